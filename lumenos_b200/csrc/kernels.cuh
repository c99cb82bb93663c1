// Kernel argument blocks and host launchers of liblumenos_b200 (definitions in kernels.cu).
#pragma once
#include <cstdint>
#include <utility>
#include <vector>
#include <cuda_runtime.h>

#include "device.cuh"

namespace lmn {

constexpr int MAX_LIMBS = 28;   // Q limbs + P limbs
constexpr int MAX_P = 4;
constexpr int MAX_DIGITS = 14;
constexpr int ENC_C = 32;       // contiguous coefficients per Encode tile row (256 B)
constexpr int ENC_MAX_TILE = 256;

// constants of the hybrid key-switch at one level (built on the host, resident in HBM)
struct KsConsts {
    int nq, np, beta, k;                       // Q limbs in use, P limbs, digits, total Q limbs (evk / col stride)
    int dig_start[MAX_DIGITS], dig_size[MAX_DIGITS];
    u64 intt_w[MAX_LIMBS], intt_ws[MAX_LIMBS];               // N^-1 (Q_d/q_i)^-1 mod q_i (Shoup pair)
    // base conversion digit d -> target tix:  V = K + sum_a z_a * c_a  (a 128-bit integer), r = V * 2^-64 mod m (Montgomery
    // step), which equals 2^(e-64) * ModUp value.  The factor 2^(64-e) is folded into the evaluation-key rows of (d, tix) when
    // the key is loaded.  For Lattigo's primes (2^b +- j*2N + 1) the constants 2^e (Q_d/q_a) mod m are tiny in signed form
    // (|c| < 2^31): "fast" pairs use 64x32-bit products.
    int mu_fast[MAX_DIGITS][MAX_LIMBS];                      // 1: small signed constants
    uint32_t mu_mag[MAX_DIGITS][MAX_LIMBS][MAX_P];           // |c_a|          (fast)
    uint32_t mu_sgn[MAX_DIGITS][MAX_LIMBS];                  // bit a: c_a < 0 (fast)
    u64 mu_c[MAX_DIGITS][MAX_LIMBS][MAX_P];                  // Q_d/q_a mod m  (generic, e = 0)
    u64 mu_k_lo[MAX_DIGITS][MAX_LIMBS], mu_k_hi[MAX_DIGITS][MAX_LIMBS];   // K = -2^e (nd-1) Q_d mod m, plus a multiple of m covering the negative terms
    u64 mu_g_w[MAX_DIGITS][MAX_LIMBS], mu_g_ws[MAX_DIGITS][MAX_LIMBS];    // 2^(e-64) mod m (Shoup pair): single-prime digits
    u64 pin_w[MAX_P], pin_ws[MAX_P], pin_add[MAX_P];         // N^-1 (P/p_j)^-1 mod p_j ; half_j (P/p_j)^-1 mod p_j
    u64 md_c[MAX_LIMBS][MAX_P];                              // (P/p_a mod q_i) * 2^64 mod q_i
    u64 md_neg[MAX_LIMBS];                                   // -((np-1) P + Phalf) mod q_i
    u64 pinv_w[MAX_LIMBS], pinv_ws[MAX_LIMBS];               // P^-1 mod q_i
    u64 pmod_w[MAX_LIMBS], pmod_ws[MAX_LIMBS];               // P mod q_i (block pipeline: "+ c0" enters as c0 P before ModDown)
};
// target index tix in [0, nq+np): modulus index in the global table
__host__ __device__ inline int ks_mod_index(int tix, int nq, int k) { return tix < nq ? tix : k + (tix - nq); }

struct RescaleConsts {   // DivRoundByLastModulusNTT from level l (limb l dropped)
    u64 half;                        // (q_l - 1) / 2
    u64 half_mod[MAX_LIMBS];         // half mod q_i
    u64 qlinv_w[MAX_LIMBS], qlinv_ws[MAX_LIMBS];
};

// the whole Rescale chain from level `from` down to level `to` (repeated DivRoundByLastModulusNTT, fhe/ligero.go:331-333,
// :149-154, :271-273), evaluated in the coefficient domain: step l (limb l dropped) uses row l
struct RescaleChain {
    int from, to;
    u64 half[MAX_LIMBS];                               // (q_l - 1) / 2
    u64 half_mod[MAX_LIMBS][MAX_LIMBS];                // [l][i] half_l mod q_i
    u64 qlinv_w[MAX_LIMBS][MAX_LIMBS], qlinv_ws[MAX_LIMBS][MAX_LIMBS];   // [l][i] q_l^-1 mod q_i (Shoup pair)
    u64 cadd[MAX_LIMBS][MAX_LIMBS];                    // [l][i] (half_l mod q_i) + ceil(q_l / q_i) q_i: x_i + cadd - r_l > 0 and = x_i - (r_l - half_l) mod q_i
};

struct DeviceTables {
    int logn, n, k, np;
    const ModDev *mods;              // [k+np (+1 small-ring modulus)]
    const ulonglong2 *tw_fwd;        // [k+np][N]
    const ulonglong2 *tw_inv;        // [k+np][N]
    const ulonglong2 *twb_fwd;       // log N >= 12: [k+np][S][2048] per-block tables in shared-memory layout (ntt_blk.cuh)
    const ulonglong2 *twb_inv;
};
void launch_build_blk_tables(cudaStream_t stream, const DeviceTables &T, ulonglong2 *fwd, ulonglong2 *inv);

struct EncPassArgs {
    const ModDev *mods;
    const u64 *in;          // slot- or position-indexed input cts
    const u64 *zero;        // ciphertext used for slots >= in_count (nullptr: none)
    u64 *out;               // position-indexed output cts
    uint32_t in_count;      // slots >= in_count read `zero`
    int in_by_pos;          // 1: input ct of slot s is in[out_pos[s]], 0: in[s]
    const uint16_t *scale_after;  // [logn][n]
    const uint32_t *out_pos;      // [n]
    const ulonglong2 *scal;       // [k][nscal]
    uint32_t nscal, n;
    int log_tile, s0, s1;   // tile = 2^log_tile slots, levels [s0, s1)
    uint32_t tile_mul, stride;    // slot = tile*tile_mul + local*stride
    uint64_t lbits;         // 4 bits per level: local index bit of level s0+i
    int k;                  // Q limbs per poly
    int logn;               // log2 of the coefficients per limb in these buffers (ring degree, or a slice of it)
    uint64_t ct_words;      // words per ciphertext (2*k*N)
    uint64_t lane_off, lane_words;   // restrict the pass to words [lane_off, lane_off + lane_words) of every ciphertext (0, 0 = all)
};

// kernel classes (profiling / launch accounting)
enum KernelId { KID_ENCODE = 0, KID_MULPT, KID_INTT_C1, KID_VFIX, KID_MAC, KID_PINTT, KID_VPFIX, KID_DOWN, KID_RS_INTT,
                KID_RS_PRE, KID_RS_APPLY, KID_PACK, KID_GATHER, KID_DECIMATE, KID_LEAF_SER, KID_SHA_LEAF, KID_SHA_NODE, KID_OTHER, KID_COUNT };
const char *kernel_name(int id);

// optional per-class device timing: CUDA event pairs recorded on the launch stream around each launch
struct Profiler {
    uint32_t mask = 0;                                    // bit i: time kernels of class i
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> spans[KID_COUNT];
    std::vector<cudaEvent_t> pool;
    uint64_t launches[KID_COUNT] = {0};
    uint64_t jobs[KID_COUNT] = {0};                       // column-jobs (ciphertexts) those launches processed
    cudaEvent_t get();
};

struct Launch {
    cudaStream_t stream;
    uint64_t *counter;      // kernel launch counter
    Profiler *prof;
};
struct LaunchScope {        // RAII: counts the launch and brackets it with events when its class is being profiled
    const Launch &L; int id; cudaEvent_t e1 = nullptr;
    LaunchScope(const Launch &l, int kid, uint64_t jobs = 0);
    ~LaunchScope();
};
// peak rate of the NTT butterfly instruction mix on the integer pipes, register-resident (no memory traffic)
double measure_peak_butterflies(cudaStream_t stream, const ModDev *mods, int sm_count, int product = 0);

// ---- elementwise / operator level ----
void launch_addsub(const Launch &L, const DeviceTables &T, int level, const u64 *a, const u64 *b, u64 *o, int sub);
void launch_mul_scalar(const Launch &L, const DeviceTables &T, int level, const u64 *a, const ulonglong2 *w_per_limb, u64 *o);
// ptm[i][n] = (pt * T mod q_i) * 2^64 mod q_i
void launch_prep_pt(const Launch &L, const DeviceTables &T, int level, const u64 *pt, u64 t_mod, u64 *ptm);
// out_x[job] = in[col] * ptm_x  (x = 0,1; ptm1 may be null); in: cts with stride in_words; out: job cols with stride col_words
void launch_mul_pt(const Launch &L, const DeviceTables &T, int level, const u64 *in, uint64_t in_words, int ncols,
                   const u64 *ptm0, const u64 *ptm1, u64 *out0, u64 *out1, uint64_t col_words);
// evk = in with rows (d, c, gm) times fm[d][gm] (Montgomery-form factors, 0 = copy); beta digits.  Out of place; for
// log N >= 12 the words of every 2048-block are stored in the block kernels' read order (key_slot in kernels.cu)
void launch_scale_rows(const Launch &L, const DeviceTables &T, const u64 *in, u64 *evk, const u64 *fm_dev, int beta);
void launch_limb_ntt(const Launch &L, const DeviceTables &T, int mod_idx, u64 *limb, int inverse);

// ---- Encode ----
void launch_encode_pass(const Launch &L, const EncPassArgs &A);

// ---- key switch + automorphism + accumulate, for `jobs` columns laid out with fixed strides ----
struct KsJobs {
    u64 *col;  uint64_t col_words;   // [jobs][2][k][N]
    u64 *col2;                       // second buffer of the same shape: the block path (log N >= 11) rotates out of place
                                     // (col -> col2) and swaps the two pointers of the view it was given
    u64 *y;    uint64_t y_words;     // [jobs][k][N]
    u64 *acc;  uint64_t acc_words;   // [jobs][2][k+np][N]
    u64 *mu;   uint64_t mu_words;    // [jobs][beta_max][k+np][N] base-converted digits (first NTT level applied when N = 16384)
    int jobs;
};
// col += pi_galEl( keyswitch(col) )      (AutomorphismHoisted + Add, lattigo rlwe/inner_sum.go)
// On return J.col holds the result (the block path has swapped J.col and J.col2).
void launch_rotate_accumulate(const Launch &L, const DeviceTables &T, const KsConsts *ksc_dev, const KsConsts &ksc_host,
                              KsJobs &J, const u64 *evk, uint32_t gal_el);
// plain key switch without automorphism into (out0,out1) = (u0 + c0, u1) at level 0 (ring switch)
// one Rescale step from level l on every job's col (in place, limb l dropped)
void launch_rescale_step(const Launch &L, const DeviceTables &T, const RescaleConsts *rc_dev, int level,
                         const KsJobs &J);
// copy limbs 0..nl-1 of both polys of each job col into a packed store [jobs][2][nl][N]
void launch_pack_out(const Launch &L, const DeviceTables &T, const KsJobs &J, int nl, u64 *out);
// gather cts: col[j] = src[idx[j]] (idx on device)
void launch_gather_in(const Launch &L, const DeviceTables &T, const u64 *src, uint64_t src_words, const uint32_t *idx,
                      const KsJobs &J);
// ring switch tail: per job, limb 0 of both polys: INTT (large), decimate by N/Ns, NTT (small) -> out [jobs][2][Ns]
void launch_ring_decimate(const Launch &L, const DeviceTables &T, const KsJobs &J, int logns, const ModDev *small_mod,
                          const ulonglong2 *small_tw_fwd, u64 *out);
// key switch at level 0 without automorphism: col[0][0] += u0 ; col[1][0] = u1
void launch_keyswitch_level0(const Launch &L, const DeviceTables &T, const KsConsts *ksc_dev, const KsConsts &ksc_host,
                             KsJobs &J, const u64 *evk);

// ---- 2048-point block pipeline (rings of 2048*S coefficients, S = 2, 4, 8; kernels_blk.cu) ----
inline int blk_blocks(int logn) { return logn >= 12 ? 1 << (logn - 11) : 0; }   // S, or 0: whole-limb kernels
struct BlkSlots {           // limb slots of one batched block transform
    const u64 *src; uint64_t src_job_words;
    u64 *dst;       uint64_t dst_job_words;
    int nslots;
    int mod[MAX_LIMBS];                                    // modulus index of the slot
    uint32_t src_limb[MAX_LIMBS], dst_limb[MAX_LIMBS];     // limb offsets (units of N words) inside one job
};
// dst = raw block INTT of src (no N^-1, the log2(S) coupling levels pending; values in [0, 4q))
void launch_blk_intt(const Launch &L, const DeviceTables &T, const BlkSlots &A, int jobs, int kid);
// y (raw block INTTs of c1) -> mu: coupling levels, scaling, float64 v, base conversion, forward coupling levels
void launch_modup_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J);
// acc P limbs (raw block INTTs) -> mu[c][i]: coupling, scaling + P/2, float64 v, P -> q_i conversion, forward coupling levels
void launch_down_pre(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J);
// block NTT of mu[c][i], (acc - ext) P^-1 (+c0), automorphism, col2 = col + pi(u)  (mode 1: col2 = u, no automorphism)
void launch_down_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J,
                     uint32_t gal_el, int mode);
// rescale step from `level`: acc slots c (raw block INTT of limb `level`) -> mu[c][i], then block NTT + apply in place
void launch_rs_pre(const Launch &L, const DeviceTables &T, const RescaleConsts *Rd, int level, const KsJobs &J);
void launch_rs_blk(const Launch &L, const DeviceTables &T, const RescaleConsts *Rd, int level, const KsJobs &J);
// all Rescale steps from level Ch.from to Ch.to at once: one raw block INTT of every limb, the chain of rounded divisions
// pointwise in the coefficient domain (k_rs_chain), one block NTT of the surviving limbs.  Bit-identical to the step-by-step
// evaluation (every step is exact arithmetic mod q_i and the NTT is linear), with 2 (from + to + 2) limb transforms per
// ciphertext instead of 2 sum_l (l + 1).
// src != nullptr: job j's input ciphertext ([2][k][N], max level) is read from src + j * src_job_words instead of J.col
void launch_rescale_chain(const Launch &L, const DeviceTables &T, const RescaleChain *Cd, const RescaleChain &Ch, const KsJobs &J,
                          const u64 *src = nullptr, uint64_t src_job_words = 0);
// block NTT + gadget MAC of the block pipeline (kernels_mac.cu)
void launch_mac_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J, const u64 *evk);
const char *kernels_mac_init();

const char *kernels_init();  // sets dynamic shared memory attributes; returns nullptr or error string

}  // namespace lmn
