// Plan compiler for the Encode stage: symbolic walk of nttInner (ref: fhe/ntt.go:20-281) over slot ids.
#include "plan.hpp"

#include <map>
#include <stdexcept>

#include "hostmath.hpp"

namespace lmn {
namespace {

struct Walker {
    EncodePlan &pl;
    const uint64_t *roots;
    uint64_t t;
    std::vector<uint32_t> level;  // butterflies seen so far per slot
    std::map<uint64_t, uint16_t> scalar_id;
    std::string err;

    Walker(EncodePlan &p, const uint64_t *r, uint64_t tt) : pl(p), roots(r), t(tt), level(p.n, 0) {}

    uint16_t id_of(uint64_t w) {
        auto it = scalar_id.find(w);
        if (it != scalar_id.end()) return it->second;
        if (pl.scalars.size() >= 65535) { err = "too many distinct twiddle scalars"; return 0; }
        uint16_t id = (uint16_t)pl.scalars.size();
        pl.scalars.push_back(w);
        scalar_id[w] = id;
        return id;
    }
    // (a, b) <- (a + b, a - b) on the ciphertexts in slots a, b   (fhe/ntt.go:26-34 and every "butterflies" block)
    void butterfly(uint32_t a, uint32_t b) {
        if (!err.empty()) return;
        uint32_t s = level[a];
        if (level[b] != s || s >= pl.logn) { err = "butterfly operands at different levels"; return; }
        uint32_t x = a ^ b;
        if (!h_is_pow2(x) || (a & x)) { err = "butterfly pair is not (bit clear, bit set) of one index bit"; return; }
        uint8_t bit = (uint8_t)h_log2(x);
        if (pl.stage_bit[s] == 0xff) pl.stage_bit[s] = bit;
        else if (pl.stage_bit[s] != bit) { err = "level pairs on more than one index bit"; return; }
        level[a] = level[b] = s + 1;
        pl.adds++; pl.subs++;
    }
    // slot <- slot * w   (backend.Mul(ct, uint64, ct), fhe/ntt.go:60,134,138,143,169,215,268)
    void scale(uint32_t slot, uint64_t w) {
        if (!err.empty()) return;
        uint32_t s = level[slot];
        if (s == 0) { err = "scalar multiply before the first butterfly"; return; }
        uint16_t &cell = pl.scale_after[(size_t)(s - 1) * pl.n + slot];
        if (cell != 0) { err = "two scalar multiplies between consecutive butterflies"; return; }
        cell = id_of(w);
        pl.muls++;
    }
};

// core.SqrtFactor (ref: core/math.go:25-35)
uint32_t sqrt_factor(uint32_t n) {
    int l = h_log2(n);
    return (l & 1) ? 1u << ((l - 1) / 2) : 1u << (l / 2);
}

// core.Transpose on the slice of slot ids (ref: core/math.go:38-60)
void transpose(uint32_t *m, uint32_t rows, uint32_t cols) {
    std::vector<uint32_t> c(m, m + (size_t)rows * cols);
    for (uint32_t i = 0; i < rows; i++)
        for (uint32_t j = 0; j < cols; j++) m[(size_t)j * rows + i] = c[(size_t)i * cols + j];
}

// radix-2 decimation-in-frequency networks of the hard-coded sizes, as (op, x, y) triples over positions:
//   'B' butterfly(x, y); 'M' multiply position x by twiddle #y; 'S' swap positions x, y
struct Op { char op; uint8_t x, y; };
// twiddle #0 = W(4), #1 = W(8), #2 = Pow(3, W(8))   (ref: fhe/ntt.go:60,134,138,142)
const Op kSize2[] = {{'B', 0, 1}};
const Op kSize4[] = {{'B', 0, 2}, {'B', 1, 3}, {'M', 3, 0}, {'B', 0, 1}, {'B', 2, 3}, {'S', 1, 2}};
const Op kSize8[] = {{'B', 0, 4}, {'B', 1, 5}, {'B', 2, 6}, {'B', 3, 7}, {'M', 5, 1}, {'M', 6, 0}, {'M', 7, 2},
                     {'B', 0, 2}, {'B', 1, 3}, {'M', 3, 0}, {'B', 0, 1}, {'B', 2, 3}, {'B', 4, 6}, {'B', 5, 7},
                     {'M', 7, 0}, {'B', 4, 5}, {'B', 6, 7}, {'S', 1, 4}, {'S', 3, 6}};

void run_fixed(Walker &w, uint32_t *v, uint32_t len, uint32_t size, const Op *ops, size_t nops) {
    uint64_t tw[3];
    tw[0] = w.roots[4 % w.pl.field_n];
    tw[1] = size >= 8 ? w.roots[8 % w.pl.field_n] : 0;
    tw[2] = size >= 8 ? h_mulmod(h_mulmod(tw[1], tw[1], w.t), tw[1], w.t) : 0;  // Field().Pow(3, RootForward(8)): plain BRed products
    for (uint32_t i = 0; i < len; i += size) {
        uint32_t *x = v + i;
        for (size_t o = 0; o < nops; o++) {
            const Op &op = ops[o];
            if (op.op == 'B') w.butterfly(x[op.x], x[op.y]);
            else if (op.op == 'M') w.scale(x[op.x], tw[op.y]);
            else { uint32_t tmp = x[op.x]; x[op.x] = x[op.y]; x[op.y] = tmp; }
        }
    }
}

void walk(Walker &w, uint32_t *v, uint32_t len, uint32_t size) {
    if (size <= 1 || !w.err.empty()) return;
    if (size == 2) return run_fixed(w, v, len, 2, kSize2, sizeof(kSize2) / sizeof(Op));
    if (size == 4) return run_fixed(w, v, len, 4, kSize4, sizeof(kSize4) / sizeof(Op));
    if (size == 8) return run_fixed(w, v, len, 8, kSize8, sizeof(kSize8) / sizeof(Op));
    // six-step split (ref: fhe/ntt.go:245-279).  `step` is declared once per call and carried across chunks.
    const uint32_t n1 = sqrt_factor(size), n2 = size / n1;
    const uint64_t fn = w.pl.field_n;
    uint64_t step = fn / size;
    for (uint32_t cs = 0; cs < len; cs += size) {
        uint32_t *chunk = v + cs;
        transpose(chunk, n1, n2);
        walk(w, chunk, size, n1);
        transpose(chunk, n2, n1);
        for (uint32_t i = 1; i < n1; i++) {
            step = ((uint64_t)i * step) % fn;
            uint64_t idx = step;
            for (uint32_t j = 1; j < n2; j++) {
                idx %= fn;
                w.scale(chunk[(size_t)i * n2 + j], w.roots[idx]);
                idx += step;
            }
        }
        walk(w, chunk, size, n2);
        transpose(chunk, n1, n2);
    }
}

}  // namespace

std::string compile_encode_plan(uint32_t n, uint32_t field_n, const uint64_t *roots, uint64_t t, uint32_t max_tile,
                                EncodePlan &pl) {
    if (!h_is_pow2(n) || n < 2) return "NTT size must be a power of two >= 2";
    if (!h_is_pow2(field_n) || field_n < n || field_n < 16) return "plaintext field degree must be a power of two >= max(size,16)";
    pl = EncodePlan();
    pl.n = n; pl.logn = (uint32_t)h_log2(n); pl.field_n = field_n;
    pl.stage_bit.assign(pl.logn, 0xff);
    pl.scale_after.assign((size_t)pl.logn * n, 0);
    pl.scalars.assign(1, 0);
    std::vector<uint32_t> v(n);
    for (uint32_t i = 0; i < n; i++) v[i] = i;
    Walker w(pl, roots, t);
    walk(w, v.data(), n, n);
    if (!w.err.empty()) return "plan compiler: " + w.err;
    for (uint32_t s = 0; s < n; s++)
        if (w.level[s] != pl.logn) return "plan compiler: slot did not pass through log2(n) butterfly levels";
    pl.out_pos.assign(n, 0);
    for (uint32_t i = 0; i < n; i++) pl.out_pos[v[i]] = i;  // returned slice position i holds slot v[i]
    // pass structure
    if (n <= max_tile) {
        pl.split = 0; pl.n1 = n; pl.n2 = 1;
    } else {
        pl.n1 = sqrt_factor(n); pl.n2 = n / pl.n1;
        if (pl.n1 > max_tile || pl.n2 > max_tile) return "NTT size too large for two device passes";
        pl.split = (uint32_t)h_log2(pl.n1);
        uint32_t lo = (uint32_t)h_log2(pl.n2);
        for (uint32_t s = 0; s < pl.logn; s++) {
            bool high = pl.stage_bit[s] >= lo;
            if ((s < pl.split) != high) return "plan compiler: levels do not separate into column and row passes";
        }
    }
    return "";
}

}  // namespace lmn
