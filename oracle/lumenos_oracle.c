/*
 * lumenos_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).  See lumenos_oracle.h.
 *
 * Every function cites the reference file:line it restates ("ref:") or the Lattigo v6
 * (v6.1.2-0.20250520151126-84f6bc33cb5b, un-vendored) routine whose published algorithm it
 * restates ("lattigo:").  All residues are canonical [0,q) at function boundaries.
 */
#include "lumenos_oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef uint64_t u64;
typedef unsigned __int128 u128;
typedef __int128 i128;

#define MAXK 24
#define MAXP 4

/* ------------------------------------------------------------------------------------------ */
/* scalar modular arithmetic                                                                  */
/* ------------------------------------------------------------------------------------------ */
static inline u64 mulmod(u64 a, u64 b, u64 m) { return (u64)(((u128)a * b) % m); }
static inline u64 addmod(u64 a, u64 b, u64 m) { u64 r = a + b; return r >= m ? r - m : r; }
static inline u64 submod(u64 a, u64 b, u64 m) { return a >= b ? a - b : a + m - b; }
static u64 powmod(u64 a, u64 e, u64 m) {
    u64 r = 1 % m; a %= m;
    while (e) { if (e & 1) r = mulmod(r, a, m); a = mulmod(a, a, m); e >>= 1; }
    return r;
}
static u64 invmod(u64 a, u64 m) { return powmod(a % m, m - 2, m); } /* m prime */
static inline u64 bitrev(u64 x, int bits) {
    u64 r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}
/* Shoup constant floor(w*2^64/q) */
static inline u64 shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
/* x*w mod q with precomputed Shoup constant; x < 2^64, result canonical */
static inline u64 mulshoup(u64 x, u64 w, u64 ws, u64 q) {
    u64 hi = (u64)(((u128)x * ws) >> 64);
    u64 r = x * w - hi * q;
    return r >= q ? r - q : r;
}

static int is_prime(u64 n) {
    if (n < 2) return 0;
    static const u64 small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (int i = 0; i < 12; i++) { if (n % small[i] == 0) return n == small[i]; }
    u64 d = n - 1; int s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (int i = 0; i < 12; i++) {
        u64 x = powmod(small[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) { x = mulmod(x, x, n); if (x == n - 1) { comp = 0; break; } }
        if (comp) return 0;
    }
    return 1;
}
static u64 gcd64(u64 a, u64 b) { while (b) { u64 t = a % b; a = b; b = t; } return a; }
static u64 pollard_rho(u64 n) {
    if (!(n & 1)) return 2;
    for (u64 c = 1;; c++) {
        u64 x = 2, y = 2, d = 1;
        while (d == 1) {
            x = addmod(mulmod(x, x, n), c, n);
            y = addmod(mulmod(y, y, n), c, n);
            y = addmod(mulmod(y, y, n), c, n);
            d = gcd64(x > y ? x - y : y - x, n);
        }
        if (d != n) return d;
    }
}
static void factor_rec(u64 n, u64 *f, int *nf) {
    if (n == 1) return;
    if (is_prime(n)) {
        for (int i = 0; i < *nf; i++) if (f[i] == n) return;
        f[(*nf)++] = n; return;
    }
    u64 d = pollard_rho(n);
    factor_rec(d, f, nf); factor_rec(n / d, f, nf);
}
/* lattigo: ring.PrimitiveRoot -- smallest generator g found by incrementing from `start` */
static u64 primitive_root(u64 q, int start) {
    u64 f[64]; int nf = 0;
    factor_rec(q - 1, f, &nf);
    for (u64 g = (u64)start;; g++) {
        int ok = 1;
        for (int i = 0; i < nf; i++) if (powmod(g, (q - 1) / f[i], q) == 1) { ok = 0; break; }
        if (ok) return g;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* per-modulus NTT tables (lattigo: ring.SubRing.generateNTTConstants, ring/ntt.go)           */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    u64 q, psi;
    int logN, N;
    u64 *rf, *rfs; /* forward roots psi^bitrev(j) (plain) + Shoup */
    u64 *rb, *rbs; /* backward roots psi^-bitrev(j) */
    u64 ninv, ninvs;
} orc_mod;

static void mod_init(orc_mod *m, u64 q, int logN, int primroot_start) {
    m->q = q; m->logN = logN; m->N = 1 << logN;
    int N = m->N;
    u64 g = primitive_root(q, primroot_start);
    m->psi = powmod(g, (q - 1) / (2 * (u64)N), q);
    u64 psiinv = invmod(m->psi, q);
    m->rf = malloc(sizeof(u64) * N); m->rfs = malloc(sizeof(u64) * N);
    m->rb = malloc(sizeof(u64) * N); m->rbs = malloc(sizeof(u64) * N);
    u64 a = 1, b = 1;
    for (int j = 0; j < N; j++) {
        int r = (int)bitrev((u64)j, logN);
        m->rf[r] = a; m->rfs[r] = shoup(a, q);
        m->rb[r] = b; m->rbs[r] = shoup(b, q);
        a = mulmod(a, m->psi, q); b = mulmod(b, psiinv, q);
    }
    m->ninv = invmod((u64)N, q); m->ninvs = shoup(m->ninv, q);
}
static void mod_free(orc_mod *m) { free(m->rf); free(m->rfs); free(m->rb); free(m->rbs); }

/* lattigo: ring.NTTStandard -- Cooley-Tukey, bit-reversed output, canonical result */
static void ntt_fwd(const orc_mod *m, u64 *a) {
    const u64 q = m->q; const int N = m->N;
    int t = N;
    for (int mm = 1; mm < N; mm <<= 1) {
        t >>= 1;
        for (int i = 0; i < mm; i++) {
            u64 w = m->rf[mm + i], ws = m->rfs[mm + i];
            u64 *x = a + 2 * i * t, *y = x + t;
            for (int j = 0; j < t; j++) {
                u64 u = x[j], v = mulshoup(y[j], w, ws, q);
                x[j] = addmod(u, v, q); y[j] = submod(u, v, q);
            }
        }
    }
}
/* lattigo: ring.INTTStandard -- Gentleman-Sande, scaled by N^-1, canonical result */
static void ntt_inv(const orc_mod *m, u64 *a) {
    const u64 q = m->q; const int N = m->N;
    int t = 1;
    for (int mm = N; mm > 1; mm >>= 1) {
        int h = mm >> 1;
        for (int i = 0; i < h; i++) {
            u64 w = m->rb[h + i], ws = m->rbs[h + i];
            u64 *x = a + 2 * i * t, *y = x + t;
            for (int j = 0; j < t; j++) {
                u64 u = x[j], v = y[j];
                x[j] = addmod(u, v, q); y[j] = mulshoup(submod(u, v, q), w, ws, q);
            }
        }
        t <<= 1;
    }
    for (int j = 0; j < N; j++) a[j] = mulshoup(a[j], m->ninv, m->ninvs, q);
}

/* ------------------------------------------------------------------------------------------ */
/* parameters                                                                                 */
/* ------------------------------------------------------------------------------------------ */
struct orc_params {
    int logN, N, k, np;
    u64 t;
    orc_mod mods[MAXK + MAXP]; /* Q primes then P primes */
    orc_mod modt;              /* ringT: degree N, modulus t (bgv.Encoder) */
    u64 *perm;                 /* bgv encoder slot permutation (lattigo: bgv.permuteMatrix) */
    orc_switches sw;
};

void orc_default_switches(orc_switches *sw) {
    sw->primroot_start = 3; sw->scalar_centered = 1; sw->digit_centered = 0;
    sw->moddown_round_nearest = 1; sw->innersum_rowswap_last = 1;
}

/* ref: fhe/bfv.go:141-158 */
int orc_num_q_primes_for_ntt(int nttSize, u64 t) {
    int bitsT = 64 - __builtin_clzll(t);
    int buffer = bitsT > 45 ? 0 : -2;
    return __builtin_ctz((unsigned)nttSize) + buffer;
}

/* lattigo: ring.NTTFriendlyPrimesGenerator.NextAlternatingPrime (state machine restated) */
typedef struct { u64 next, prev, nth; int size, cn, cp; } primegen;
static void primegen_init(primegen *g, int bits, u64 nth) {
    g->size = bits; g->nth = nth; g->cn = 1; g->cp = 1;
    g->next = (1ULL << bits) + 1; g->prev = (1ULL << bits) + 1;
    if (g->prev < nth) g->cp = 0;
    g->prev -= nth;
}
static u64 primegen_next(primegen *g) {
    for (;;) {
        if (!(g->cn || g->cp)) { fprintf(stderr, "oracle: prime generator exhausted\n"); abort(); }
        if (g->cn) {
            if ((64 - __builtin_clzll(g->next)) > g->size + 1 || g->next > 0xffffffffffffffffULL - g->nth) {
                g->cn = 0;
            } else {
                u64 c = g->next; g->next += g->nth;
                if (is_prime(c)) return c;
            }
        }
        if (g->cp) {
            if ((64 - __builtin_clzll(g->prev)) < g->size || g->prev < g->nth) {
                g->cp = 0;
            } else {
                u64 c = g->prev; g->prev -= g->nth;
                if (is_prime(c)) return c;
            }
        }
    }
}

static orc_params *params_build(int logN, int k, const u64 *q, int np, const u64 *p, u64 t,
                                const orc_switches *sw) {
    if (k > MAXK || np > MAXP || k < 1) return NULL;
    orc_params *P = calloc(1, sizeof(*P));
    P->logN = logN; P->N = 1 << logN; P->k = k; P->np = np; P->t = t;
    if (sw) P->sw = *sw; else orc_default_switches(&P->sw);
    for (int i = 0; i < k; i++) mod_init(&P->mods[i], q[i], logN, P->sw.primroot_start);
    for (int j = 0; j < np; j++) mod_init(&P->mods[k + j], p[j], logN, P->sw.primroot_start);
    mod_init(&P->modt, t, logN, P->sw.primroot_start);
    /* lattigo: bgv.permuteMatrix */
    int N = P->N;
    P->perm = malloc(sizeof(u64) * N);
    u64 pow = 1, mask = 2 * (u64)N - 1;
    for (int i = 0, j = N / 2; i < N / 2; i++, j++) {
        u64 pos = bitrev(pow >> 1, logN);
        P->perm[i] = pos; P->perm[j] = (u64)N - pos - 1;
        pow = (pow * 5) & mask;
    }
    return P;
}

/* ref: fhe/bfv.go:160-185 (LogQ=[58,56,...], LogP=[55,55]); lattigo: rlwe.GenModuli */
orc_params *orc_params_generate(int logN, int k, u64 t, const orc_switches *sw) {
    u64 nth = 2ULL << logN;
    if (t % nth != 1) return NULL; /* ref: fhe/bfv.go:136-140 */
    u64 q[MAXK], p[2];
    primegen g58, g56, g55;
    primegen_init(&g58, 58, nth); primegen_init(&g56, 56, nth); primegen_init(&g55, 55, nth);
    q[0] = primegen_next(&g58);
    for (int i = 1; i < k; i++) q[i] = primegen_next(&g56);
    p[0] = primegen_next(&g55); p[1] = primegen_next(&g55);
    return params_build(logN, k, q, 2, p, t, sw);
}
orc_params *orc_params_from_moduli(int logN, int k, const u64 *q, int np, const u64 *p, u64 t,
                                   const orc_switches *sw) {
    return params_build(logN, k, q, np, p, t, sw);
}
void orc_params_free(orc_params *P) {
    if (!P) return;
    for (int i = 0; i < P->k + P->np; i++) mod_free(&P->mods[i]);
    mod_free(&P->modt); free(P->perm); free(P);
}
int orc_params_k(const orc_params *P) { return P->k; }
int orc_params_np(const orc_params *P) { return P->np; }
int orc_params_N(const orc_params *P) { return P->N; }
u64 orc_params_t(const orc_params *P) { return P->t; }
void orc_params_moduli(const orc_params *P, u64 *q, u64 *p) {
    for (int i = 0; i < P->k; i++) q[i] = P->mods[i].q;
    for (int j = 0; j < P->np; j++) p[j] = P->mods[P->k + j].q;
}
void orc_params_psi(const orc_params *P, u64 *psi_q, u64 *psi_p) {
    for (int i = 0; i < P->k; i++) psi_q[i] = P->mods[i].psi;
    for (int j = 0; j < P->np; j++) psi_p[j] = P->mods[P->k + j].psi;
}
u64 orc_params_psi_t(const orc_params *P) { return P->modt.psi; }   /* 2N-th root of the encoder's ring T (lattigo: ringT) */
int orc_beta(const orc_params *P) { return (P->k + P->np - 1) / P->np; }

void orc_ntt_limb(const orc_params *P, int mi, u64 *a) { ntt_fwd(&P->mods[mi], a); }
void orc_intt_limb(const orc_params *P, int mi, u64 *a) { ntt_inv(&P->mods[mi], a); }

/* lattigo: ring.AutomorphismNTTIndex */
void orc_automorphism_index(const orc_params *P, u64 galEl, uint32_t *index) {
    int logN = P->logN; u64 mask = 2 * (u64)P->N - 1;
    for (int i = 0; i < P->N; i++) {
        u64 t1 = 2 * bitrev((u64)i, logN) + 1;
        u64 t2 = (((galEl * t1) & mask) - 1) >> 1;
        index[i] = (uint32_t)bitrev(t2, logN);
    }
}

/* ------------------------------------------------------------------------------------------ */
/* witness generator: ChaCha20 (IETF) keystream, ref: core/utils.go:46-82                     */
/* ------------------------------------------------------------------------------------------ */
#define ROTL32(v, n) (((v) << (n)) | ((v) >> (32 - (n))))
#define QR(a, b, c, d) a += b; d ^= a; d = ROTL32(d, 16); c += d; b ^= c; b = ROTL32(b, 12); \
                       a += b; d ^= a; d = ROTL32(d, 8);  c += d; b ^= c; b = ROTL32(b, 7);
static void chacha20_block(const uint32_t in[16], uint8_t out[64]) {
    uint32_t x[16];
    memcpy(x, in, 64);
    for (int i = 0; i < 10; i++) {
        QR(x[0], x[4], x[8], x[12]) QR(x[1], x[5], x[9], x[13]) QR(x[2], x[6], x[10], x[14]) QR(x[3], x[7], x[11], x[15])
        QR(x[0], x[5], x[10], x[15]) QR(x[1], x[6], x[11], x[12]) QR(x[2], x[7], x[8], x[13]) QR(x[3], x[4], x[9], x[14])
    }
    for (int i = 0; i < 16; i++) {
        uint32_t v = x[i] + in[i];
        out[4 * i] = (uint8_t)v; out[4 * i + 1] = (uint8_t)(v >> 8);
        out[4 * i + 2] = (uint8_t)(v >> 16); out[4 * i + 3] = (uint8_t)(v >> 24);
    }
}
void orc_witness_matrix(int rows, int cols, u64 t, u64 *out) {
    /* key = LE64(1) || 0^24, nonce = 0^12, counter = 0; keystream words LE u64 % t, row-major */
    uint32_t st[16] = {0x61707865, 0x3320646e, 0x79622d32, 0x6b206574, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    uint8_t blk[64];
    size_t total = (size_t)rows * cols;
    for (size_t w = 0; w < total; w += 8) {
        chacha20_block(st, blk);
        st[12]++;
        for (int j = 0; j < 8 && w + j < total; j++) {
            u64 v = 0;
            for (int b = 7; b >= 0; b--) v = (v << 8) | blk[8 * j + b];
            out[w + j] = v % t;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* plaintext field twin: core/field.go:138-197 (tables), core/ntt.go, core/code.go, core/math.go */
/* ------------------------------------------------------------------------------------------ */
int orc_field_roots_forward(u64 t, int Nfield, int primroot_start, u64 *table) {
    if (Nfield <= 0 || (Nfield & (Nfield - 1))) return -1;
    u64 nth = 2 * (u64)Nfield;
    if (!is_prime(t) || (t & (nth - 1)) != 1) return -1; /* ref: core/field.go:147-153 */
    int logN = __builtin_ctz((unsigned)Nfield);
    u64 g = primitive_root(t, primroot_start);
    u64 psi = powmod(g, (t - 1) / nth, t);
    u64 R = (u64)((((u128)1) << 64) % t); /* MForm(1) */
    u64 cur = R;                          /* psi^j * 2^64 mod t */
    for (u64 j = 0; j < (u64)Nfield; j++) {
        table[bitrev(j, logN)] = cur;
        cur = mulmod(cur, psi, t);
    }
    return 0;
}

/* ref: core/math.go:25-35 */
static int sqrt_factor(int n) {
    int l = 31 - __builtin_clz((unsigned)n);
    return (l & 1) ? 1 << ((l - 1) / 2) : 1 << (l / 2);
}

/*
 * Generic restatement of nttInner (ref: fhe/ntt.go:20-281 and its twin core/ntt.go:9-98).
 * The element type is abstracted: `idx[]` is the slice of "pointers" (element ids) that the
 * reference permutes with core.Transpose; butterflies and scalar multiplies act on the
 * elements the ids designate.
 */
typedef struct nttctx {
    void (*bfly)(struct nttctx *, int ida, int idb);    /* (a,b) <- (a+b, a-b)   */
    void (*mulw)(struct nttctx *, int id, u64 w);        /* a <- a * w            */
    const u64 *roots; int Nfield; u64 t;
    u64 n_add, n_sub, n_mul;
    void *user;
} nttctx;

/* ref: core/math.go:38-60 (Transpose of the slice) */
static void transpose_ids(int *m, int rows, int cols) {
    if (rows == cols) {
        for (int i = 0; i < rows; i++)
            for (int j = i + 1; j < cols; j++) { int tmp = m[i * cols + j]; m[i * cols + j] = m[j * rows + i]; m[j * rows + i] = tmp; }
    } else {
        int *c = malloc(sizeof(int) * rows * cols);
        memcpy(c, m, sizeof(int) * rows * cols);
        for (int i = 0; i < rows; i++)
            for (int j = 0; j < cols; j++) m[j * rows + i] = c[i * cols + j];
        free(c);
    }
}
static inline void BF(nttctx *c, int *v, int a, int b) { c->bfly(c, v[a], v[b]); c->n_add++; c->n_sub++; }
static inline void MW(nttctx *c, int *v, int a, u64 w) { c->mulw(c, v[a], w); c->n_mul++; }
static inline void SWAP(int *v, int a, int b) { int t = v[a]; v[a] = v[b]; v[b] = t; }

static void ntt_inner(nttctx *c, int *v, int len, int size) {
    switch (size) {
    case 0: case 1: return;
    case 2: /* ref: fhe/ntt.go:24-35 */
        for (int i = 0; i < len; i += 2) BF(c, v + i, 0, 1);
        return;
    case 4: { /* ref: fhe/ntt.go:36-89 */
        u64 w4 = c->roots[4];
        for (int i = 0; i < len; i += 4) {
            int *x = v + i;
            BF(c, x, 0, 2); BF(c, x, 1, 3);
            MW(c, x, 3, w4);
            BF(c, x, 0, 1); BF(c, x, 2, 3);
            SWAP(x, 1, 2);
        }
        return;
    }
    case 8: { /* ref: fhe/ntt.go:90-244 */
        u64 w4 = c->roots[4], w8 = c->roots[8];
        u64 w8_3 = mulmod(mulmod(w8, w8, c->t), w8, c->t); /* Field().Pow(3, RootForward(8)), ref: fhe/ntt.go:142, core/field.go:101-128 */
        for (int i = 0; i < len; i += 8) {
            int *x = v + i;
            BF(c, x, 0, 4); BF(c, x, 1, 5); BF(c, x, 2, 6); BF(c, x, 3, 7);
            MW(c, x, 5, w8); MW(c, x, 6, w4); MW(c, x, 7, w8_3);
            BF(c, x, 0, 2); BF(c, x, 1, 3);
            MW(c, x, 3, w4);
            BF(c, x, 0, 1); BF(c, x, 2, 3); BF(c, x, 4, 6); BF(c, x, 5, 7);
            MW(c, x, 7, w4);
            BF(c, x, 4, 5); BF(c, x, 6, 7);
            SWAP(x, 1, 4); SWAP(x, 3, 6);
        }
        return;
    }
    default: { /* six-step, ref: fhe/ntt.go:245-279 */
        int n1 = sqrt_factor(size), n2 = size / n1;
        long long step = c->Nfield / size; /* declared once, carried across chunks */
        for (int cs = 0; cs < len; cs += size) {
            int *chunk = v + cs;
            transpose_ids(chunk, n1, n2);
            ntt_inner(c, chunk, size, n1);
            transpose_ids(chunk, n2, n1);
            for (int i = 1; i < n1; i++) {
                step = ((long long)i * step) % c->Nfield;
                long long idx = step;
                for (int j = 1; j < n2; j++) {
                    idx %= c->Nfield;
                    MW(c, chunk, i * n2 + j, c->roots[idx]);
                    idx += step;
                }
            }
            ntt_inner(c, chunk, size, n2);
            transpose_ids(chunk, n1, n2);
        }
        return;
    }
    }
}

/* --- plaintext instantiation (core/ntt.go): field.Add/Sub = CRed, field.Mul = BRed --- */
typedef struct { u64 *data; } plainuser;
static void plain_bfly(nttctx *c, int a, int b) {
    u64 *d = ((plainuser *)c->user)->data; u64 x = d[a], y = d[b];
    d[a] = addmod(x, y, c->t); d[b] = submod(x, y, c->t);
}
static void plain_mulw(nttctx *c, int a, u64 w) {
    u64 *d = ((plainuser *)c->user)->data; d[a] = mulmod(d[a], w, c->t);
}
void orc_core_ntt(u64 *v, int len, int size, u64 t, const u64 *roots, int Nfield) {
    int *ids = malloc(sizeof(int) * len);
    for (int i = 0; i < len; i++) ids[i] = i;
    plainuser pu = {v};
    nttctx c = {plain_bfly, plain_mulw, roots, Nfield, t, 0, 0, 0, &pu};
    ntt_inner(&c, ids, len, size);
    u64 *tmp = malloc(sizeof(u64) * len);
    for (int i = 0; i < len; i++) tmp[i] = v[ids[i]];
    memcpy(v, tmp, sizeof(u64) * len);
    free(tmp); free(ids);
}
/* ref: core/code.go:3-23; the field is core.NewPrimeField(t, cols*rhoInv) (cmd/server/main.go:57) */
void orc_core_encode(const u64 *row, int cols, int rhoInv, u64 t, const u64 *roots, u64 *out) {
    int n = cols * rhoInv;
    memcpy(out, row, sizeof(u64) * cols);
    memset(out + cols, 0, sizeof(u64) * (n - cols));
    orc_core_ntt(out, n, n, t, roots, n);
}

/* ref: core/math.go:8-21 InnerProduct over every column of a row-major matrix, as LigeroProveReference computes MatR / MatZ
 * (fhe/ligero.go:799-953): out[j] = sum_i (r_i mod t) M[i][j] mod t */
void orc_vec_mat_mod(const u64 *r, const u64 *M_rowmajor, int rows, int cols, u64 t, u64 *out) {
    memset(out, 0, sizeof(u64) * (size_t)cols);
    for (int i = 0; i < rows; i++) {
        const u64 ri = r[i] % t; const u64 *row = M_rowmajor + (size_t)i * cols;
        for (int j = 0; j < cols; j++) out[j] = addmod(out[j], mulmod(ri, row[j], t), t);
    }
}

/* ------------------------------------------------------------------------------------------ */
/* ServerBFV operator level                                                                   */
/* ------------------------------------------------------------------------------------------ */
#define LIMB(ct, level, N, p, i) ((ct) + ((size_t)(p) * ((level) + 1) + (i)) * (size_t)(N))

/* lattigo: bgv.Evaluator.Add / ring.Add (ref call sites: fhe/ntt.go:27,40,...) */
void orc_ct_add(const orc_params *P, int level, const u64 *a, const u64 *b, u64 *out) {
    int N = P->N;
    for (int p = 0; p < 2; p++) for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q; const u64 *x = LIMB(a, level, N, p, i), *y = LIMB(b, level, N, p, i);
        u64 *o = LIMB(out, level, N, p, i);
        for (int j = 0; j < N; j++) o[j] = addmod(x[j], y[j], q);
    }
}
void orc_ct_sub(const orc_params *P, int level, const u64 *a, const u64 *b, u64 *out) {
    int N = P->N;
    for (int p = 0; p < 2; p++) for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q; const u64 *x = LIMB(a, level, N, p, i), *y = LIMB(b, level, N, p, i);
        u64 *o = LIMB(out, level, N, p, i);
        for (int j = 0; j < N; j++) o[j] = submod(x[j], y[j], q);
    }
}
/* lattigo: bgv.Evaluator.Mul(ct, uint64) -> scalar mod t, centred, MulScalarBigint
   (ref: fhe/bfv.go:34-37, call sites fhe/ntt.go:60,134,138,143,169,215,268) */
static u64 scalar_mod_qi(const orc_params *P, u64 w, u64 qi) {
    u64 t = P->t; w %= t;
    if (P->sw.scalar_centered && w > (t >> 1)) { /* w - t, Euclidean mod qi */
        u64 neg = (t - w) % qi;
        return neg ? qi - neg : 0;
    }
    return w % qi;
}
void orc_ct_mul_scalar(const orc_params *P, int level, const u64 *a, u64 w, u64 *out) {
    int N = P->N;
    for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q, wi = scalar_mod_qi(P, w, q), ws = shoup(wi, q);
        for (int p = 0; p < 2; p++) {
            const u64 *x = LIMB(a, level, N, p, i); u64 *o = LIMB(out, level, N, p, i);
            for (int j = 0; j < N; j++) o[j] = mulshoup(x[j], wi, ws, q);
        }
    }
}
/* lattigo: bgv.Evaluator.MulNew(ct, pt) -> tensorStandard, degree-0 operand:
   out_p = ct_p * pt * (T mod q_i)   (ref: fhe/bfv.go:39-42, fhe/ligero.go:319) */
void orc_ct_mul_pt(const orc_params *P, int level, const u64 *a, const u64 *pt, u64 *out) {
    int N = P->N;
    for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q, tq = P->t % q;
        const u64 *m = pt + (size_t)i * N;
        for (int p = 0; p < 2; p++) {
            const u64 *x = LIMB(a, level, N, p, i); u64 *o = LIMB(out, level, N, p, i);
            for (int j = 0; j < N; j++) o[j] = mulmod(mulmod(m[j], tq, q), x[j], q);
        }
    }
}

/* ---- hybrid key-switch (lattigo: rlwe.Evaluator.DecomposeNTT / GadgetProduct / ModDown) ---- */

/* product of the digit's primes present at `level`, reduced mod m, excluding index `skip` (-1: none) */
static u64 digit_prod_mod(const orc_params *P, int d, int level, int skip, u64 m) {
    u64 r = 1 % m;
    int s = d * P->np, e = s + P->np; if (e > level + 1) e = level + 1;
    for (int i = s; i < e; i++) if (i != skip) r = mulmod(r, P->mods[i].q % m, m);
    return r;
}

/* lattigo: ring.Decomposer.DecomposeAndSplit + rlwe DecomposeSingleNTT.
   c1intt: [level+1][N] coefficient domain; c1ntt: NTT domain; out: [k+np][N] NTT domain
   (rows for Q limbs > level are left untouched). */
static void decompose_digit(const orc_params *P, int level, int d, const u64 *c1ntt, const u64 *c1intt, u64 *out) {
    int N = P->N, k = P->k, np = P->np;
    int s = d * np, e = s + np; if (e > level + 1) e = level + 1;
    int nd = e - s;
    /* target moduli: Q limbs 0..level (not in digit) and all P limbs */
    for (int tm = 0; tm < k + np; tm++) {
        if (tm < k && tm > level) continue;
        u64 *o = out + (size_t)tm * N;
        if (tm >= s && tm < e) { memcpy(o, c1ntt + (size_t)tm * N, sizeof(u64) * N); continue; }
        const orc_mod *M = &P->mods[tm]; u64 m = M->q;
        if (nd == 1) {
            /* single-prime digit: centred lift (DecomposeAndSplit, decompLvl < 0 branch) */
            u64 qd = P->mods[s].q; const u64 *x = c1intt + (size_t)s * N;
            for (int j = 0; j < N; j++) {
                u64 c = x[j];
                if (c >= (qd >> 1)) { u64 r = (qd - c) % m; o[j] = r ? m - r : 0; }
                else o[j] = c % m;
            }
        } else {
            /* exact base conversion with float64 overflow estimate (reconstructRNS + multSum) */
            u64 yc[MAXP], qo[MAXP], Qd = digit_prod_mod(P, d, level, -1, m);
            for (int a = 0; a < nd; a++) {
                u64 qa = P->mods[s + a].q;
                yc[a] = invmod(digit_prod_mod(P, d, level, s + a, qa), qa); /* (Q_d/q_a)^-1 mod q_a */
                qo[a] = digit_prod_mod(P, d, level, s + a, m);              /* Q_d/q_a mod m */
            }
            u64 Qhalf_m = 0; /* only used by the digit_centered hypothesis */
            for (int j = 0; j < N; j++) {
                double vf = 0.0; u128 acc = 0;
                for (int a = 0; a < nd; a++) {
                    u64 qa = P->mods[s + a].q;
                    u64 y = mulmod(c1intt[(size_t)(s + a) * N + j], yc[a], qa);
                    vf += (double)y / (double)qa;
                    acc += (u128)y * qo[a];
                }
                u64 v = (u64)vf;
                u64 r = (u64)(acc % m);
                r = submod(r, mulmod(v % m, Qd, m), m);
                if (P->sw.digit_centered) {
                    /* hypothesis: representative in (-Q_d/2, Q_d/2]; decided on the float estimate of the fraction */
                    double frac = vf - (double)v;
                    if (frac >= 0.5) r = submod(r, Qd, m);
                    (void)Qhalf_m;
                }
                o[j] = r;
            }
        }
        ntt_fwd(M, o);
    }
}

/* lattigo: ring.BasisExtender.ModDownQPtoQNTT.  accQ: [level+1][N] (stride rows of N starting at accQ),
   accP: [np][N], both NTT; out: [level+1][N] NTT */
static void moddown(const orc_params *P, int level, const u64 *accQ, const u64 *accP, u64 *out) {
    int N = P->N, k = P->k, np = P->np;
    u64 *pc = malloc(sizeof(u64) * (size_t)np * N);
    u64 *ext = malloc(sizeof(u64) * N);
    /* P product & half as 128/256-bit: np<=2 fits u128; for np>2 use per-modulus arithmetic */
    for (int j = 0; j < np; j++) {
        memcpy(pc + (size_t)j * N, accP + (size_t)j * N, sizeof(u64) * N);
        ntt_inv(&P->mods[k + j], pc + (size_t)j * N);
    }
    /* PHalf = floor(P/2) = (P-1)/2 ; residues: mod p_j = (p_j-1)/2 , mod m = (P mod m - 1)*inv2 */
    u64 yc[MAXP];
    for (int a = 0; a < np; a++) {
        u64 pa = P->mods[k + a].q, pr = 1;
        for (int b = 0; b < np; b++) if (b != a) pr = mulmod(pr, P->mods[k + b].q % pa, pa);
        yc[a] = invmod(pr, pa);
        if (P->sw.moddown_round_nearest) {
            u64 h = (pa - 1) >> 1;
            u64 *x = pc + (size_t)a * N;
            for (int j = 0; j < N; j++) x[j] = addmod(x[j], h, pa);
        }
        u64 *x = pc + (size_t)a * N;
        for (int j = 0; j < N; j++) x[j] = mulmod(x[j], yc[a], pa); /* y_a */
    }
    for (int i = 0; i <= level; i++) {
        const orc_mod *M = &P->mods[i]; u64 m = M->q;
        u64 po[MAXP], Pm = 1;
        for (int a = 0; a < np; a++) {
            Pm = mulmod(Pm, P->mods[k + a].q % m, m);
            u64 pr = 1;
            for (int b = 0; b < np; b++) if (b != a) pr = mulmod(pr, P->mods[k + b].q % m, m);
            po[a] = pr;
        }
        u64 half_m = mulmod(submod(Pm, 1, m), invmod(2, m), m);
        u64 pinv = invmod(Pm, m), pinvs = shoup(pinv, m);
        for (int j = 0; j < N; j++) {
            double vf = 0.0; u128 acc = 0;
            for (int a = 0; a < np; a++) {
                u64 y = pc[(size_t)a * N + j];
                vf += (double)y / (double)P->mods[k + a].q;
                acc += (u128)y * po[a];
            }
            u64 v = (u64)vf;
            u64 r = submod((u64)(acc % m), mulmod(v % m, Pm, m), m);
            if (P->sw.moddown_round_nearest) r = submod(r, half_m, m);
            ext[j] = r;
        }
        ntt_fwd(M, ext);
        const u64 *x = accQ + (size_t)i * N; u64 *o = out + (size_t)i * N;
        for (int j = 0; j < N; j++) o[j] = mulshoup(submod(x[j], ext[j], m), pinv, pinvs, m);
    }
    free(pc); free(ext);
}

/* Montgomery reduce helper for evk (NTT + Montgomery form): a * b_mont * 2^-64 mod q */
static inline u64 mont_mul(u64 a, u64 bm, u64 q, u64 qinv /* q^-1 mod 2^64 */) {
    u128 m = (u128)a * bm;
    u64 mlo = (u64)m, mhi = (u64)(m >> 64);
    u64 hhi = (u64)(((u128)(mlo * qinv) * q) >> 64);
    u64 r = mhi - hhi + q;
    return r >= q ? r - q : r;
}
static u64 qinv64(u64 q) { u64 x = 1; for (int i = 0; i < 63; i++) { x *= q; q *= q; } return x; } /* lattigo: ring.GenMRedConstant */

/* lattigo: GadgetProduct (= gadgetProductMultiplePLazy + ModDown): cx [level+1][N] NTT;
   evk [beta_max][2][k+np][N]; out0,out1 [level+1][N] */
static void gadget_product(const orc_params *P, int level, const u64 *cx, const u64 *evk, u64 *out0, u64 *out1) {
    int N = P->N, k = P->k, np = P->np, KP = k + np;
    int beta = (level + 1 + np - 1) / np;
    u64 *cinv = malloc(sizeof(u64) * (size_t)(level + 1) * N);
    u64 *dec = malloc(sizeof(u64) * (size_t)KP * N);
    u64 *acc0 = calloc((size_t)KP * N, sizeof(u64)), *acc1 = calloc((size_t)KP * N, sizeof(u64));
    memcpy(cinv, cx, sizeof(u64) * (size_t)(level + 1) * N);
    for (int i = 0; i <= level; i++) ntt_inv(&P->mods[i], cinv + (size_t)i * N);
    for (int d = 0; d < beta; d++) {
        decompose_digit(P, level, d, cx, cinv, dec);
        for (int tm = 0; tm < KP; tm++) {
            if (tm < k && tm > level) continue;
            u64 q = P->mods[tm].q, qi = qinv64(q);
            const u64 *e0 = evk + (((size_t)d * 2 + 0) * KP + tm) * N, *e1 = evk + (((size_t)d * 2 + 1) * KP + tm) * N;
            const u64 *x = dec + (size_t)tm * N;
            u64 *a0 = acc0 + (size_t)tm * N, *a1 = acc1 + (size_t)tm * N;
            for (int j = 0; j < N; j++) {
                a0[j] = addmod(a0[j], mont_mul(x[j] % q, e0[j], q, qi), q);
                a1[j] = addmod(a1[j], mont_mul(x[j] % q, e1[j], q, qi), q);
            }
        }
    }
    moddown(P, level, acc0, acc0 + (size_t)k * N, out0);
    moddown(P, level, acc1, acc1 + (size_t)k * N, out1);
    free(cinv); free(dec); free(acc0); free(acc1);
}

/* lattigo: rlwe.Evaluator.Automorphism / AutomorphismHoisted: out = pi_galEl( (u0 + c0, u1) ) */
void orc_ct_automorphism(const orc_params *P, int level, const u64 *in, u64 galEl, const u64 *evk, u64 *out) {
    int N = P->N; size_t L = (size_t)(level + 1) * N;
    if (galEl == 1) { memmove(out, in, sizeof(u64) * 2 * L); return; }
    u64 *u0 = malloc(sizeof(u64) * L), *u1 = malloc(sizeof(u64) * L);
    uint32_t *idx = malloc(sizeof(uint32_t) * N);
    gadget_product(P, level, in + L, evk, u0, u1);
    for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q;
        for (int j = 0; j < N; j++) u0[(size_t)i * N + j] = addmod(u0[(size_t)i * N + j], in[(size_t)i * N + j], q);
    }
    orc_automorphism_index(P, galEl, idx);
    for (int i = 0; i <= level; i++)
        for (int j = 0; j < N; j++) {
            out[(size_t)i * N + j] = u0[(size_t)i * N + idx[j]];
            out[L + (size_t)i * N + j] = u1[(size_t)i * N + idx[j]];
        }
    free(u0); free(u1); free(idx);
}

static u64 galois_element(const orc_params *P, int rot) { /* lattigo: Parameters.GaloisElement(k) = 5^k mod 2N */
    u64 mask = 2 * (u64)P->N - 1, g = 1, b = 5; unsigned e = (unsigned)rot & (unsigned)(P->N / 2 - 1);
    /* 5 has order N/2 modulo 2N */
    while (e) { if (e & 1) g = (g * b) & mask; b = (b * b) & mask; e >>= 1; }
    return g;
}

/* lattigo: bgv.Parameters.GaloisElementsForInnerSum(1, n) (ref: fhe/ligero_test.go:53) */
int orc_galois_elements_for_innersum(const orc_params *P, int n, u64 *galEls) {
    int cnt = 0, half = P->N / 2;
    int m = n < half ? n : half;
    for (int s = 1; s < m; s <<= 1) galEls[cnt++] = galois_element(P, s);
    if (n == P->N) galEls[cnt++] = 2 * (u64)P->N - 1; /* row swap */
    return cnt;
}

static const u64 *find_key(u64 galEl, const u64 *galEls, const u64 *const *evks, int nkeys) {
    for (int i = 0; i < nkeys; i++) if (galEls[i] == galEl) return evks[i];
    return NULL;
}

/* lattigo: rlwe.Evaluator.InnerSum(ct, 1, n) for power-of-two n, plus the bgv-level row-swap step
   when n == N.  (ref: fhe/ligero.go:325) */
int orc_ct_innersum(const orc_params *P, int level, u64 *ct, int n, const u64 *galEls,
                    const u64 *const *evks, int nkeys) {
    int N = P->N; size_t L2 = 2 * (size_t)(level + 1) * N;
    if (n <= 0 || (n & (n - 1)) || n > N) return -1;
    u64 *rot = malloc(sizeof(u64) * L2);
    int full = (n == N), m = full ? N / 2 : n;
    int rc = 0;
    /* order: [row swap first?] -> log2(m) column rotations -> [row swap last?] */
    for (int phase = 0; phase < 3 && rc == 0; phase++) {
        int swap_phase = P->sw.innersum_rowswap_last ? 2 : 0;
        if (phase == 1) {
            for (int s = 1; s < m; s <<= 1) {
                u64 g = galois_element(P, s);
                const u64 *key = find_key(g, galEls, evks, nkeys);
                if (!key) { rc = -2; break; }
                orc_ct_automorphism(P, level, ct, g, key, rot);
                orc_ct_add(P, level, ct, rot, ct);
            }
        } else if (full && phase == swap_phase) {
            const u64 *key = find_key(2 * (u64)N - 1, galEls, evks, nkeys);
            if (!key) { rc = -2; break; }
            orc_ct_automorphism(P, level, ct, 2 * (u64)N - 1, key, rot);
            orc_ct_add(P, level, ct, rot, ct);
        }
    }
    free(rot);
    return rc;
}

/* lattigo: bgv.Evaluator.Rescale -> ring.DivRoundByLastModulusNTT (ref: fhe/ligero.go:331-333,149-154,271-273) */
u64 orc_ct_rescale(const orc_params *P, int level, const u64 *in, u64 *out, u64 scale) {
    int N = P->N;
    u64 ql = P->mods[level].q, half = (ql - 1) >> 1;
    u64 *r = malloc(sizeof(u64) * N), *tmp = malloc(sizeof(u64) * N);
    for (int p = 0; p < 2; p++) {
        memcpy(r, LIMB(in, level, N, p, level), sizeof(u64) * N);
        ntt_inv(&P->mods[level], r);
        for (int j = 0; j < N; j++) r[j] = addmod(r[j], half, ql);
        for (int i = 0; i < level; i++) {
            const orc_mod *M = &P->mods[i]; u64 m = M->q;
            u64 hm = half % m, qlinv = invmod(ql % m, m), qs = shoup(qlinv, m);
            for (int j = 0; j < N; j++) tmp[j] = submod(r[j] % m, hm, m);
            ntt_fwd(M, tmp);
            const u64 *x = LIMB(in, level, N, p, i); u64 *o = LIMB(out, level - 1, N, p, i);
            for (int j = 0; j < N; j++) o[j] = mulshoup(submod(x[j], tmp[j], m), qlinv, qs, m);
        }
    }
    free(r); free(tmp);
    /* Scale <- Scale / q_l mod t */
    return mulmod(scale % P->t, invmod(ql % P->t, P->t), P->t);
}

u64 orc_rescale_to_level1(const orc_params *P, const u64 *ct, u64 *out) {
    int N = P->N, level = P->k - 1; u64 scale = 1;
    if (level <= 1) { memcpy(out, ct, sizeof(u64) * 2 * (size_t)(level + 1) * N); return scale; }
    u64 *a = malloc(sizeof(u64) * 2 * (size_t)(level + 1) * N), *b = malloc(sizeof(u64) * 2 * (size_t)(level + 1) * N);
    memcpy(a, ct, sizeof(u64) * 2 * (size_t)(level + 1) * N);
    while (level > 1) { scale = orc_ct_rescale(P, level, a, b, scale); level--; u64 *t = a; a = b; b = t; }
    memcpy(out, a, sizeof(u64) * 2 * 2 * (size_t)N);
    free(a); free(b);
    return scale;
}

/* ref: fhe/ligero.go:318-334 (one iteration of the worker loop, ring switch excluded) */
u64 orc_inner_sum_eval_column(const orc_params *P, const u64 *ct, const u64 *pt, int rows,
                              const u64 *galEls, const u64 *const *evks, int nkeys, u64 *out) {
    int N = P->N, level = P->k - 1;
    u64 *col = malloc(sizeof(u64) * 2 * (size_t)(level + 1) * N);
    orc_ct_mul_pt(P, level, ct, pt, col);
    if (orc_ct_innersum(P, level, col, rows, galEls, evks, nkeys) != 0) { free(col); return 0; }
    u64 scale = orc_rescale_to_level1(P, col, out);
    free(col);
    return scale;
}

/* lattigo: rlwe.Evaluator.ApplyEvaluationKey (large -> small ring), ref: fhe/ring_switch.go:106-113.
   Key switch at level min(1, 0) = 0, then per limb INTT, decimate, NTT in the small ring. */
void orc_ct_ringswitch(const orc_params *P, const u64 *in, const u64 *evk, int logNs, u64 *out) {
    int N = P->N, Ns = 1 << logNs, gap = N / Ns;
    /* level-1 input [2][2][N]; only limb 0 takes part */
    u64 *c0 = malloc(sizeof(u64) * N), *c1 = malloc(sizeof(u64) * N), *u0 = malloc(sizeof(u64) * N), *u1 = malloc(sizeof(u64) * N);
    memcpy(c0, in, sizeof(u64) * N); memcpy(c1, in + 2 * (size_t)N, sizeof(u64) * N);
    gadget_product(P, 0, c1, evk, u0, u1);
    u64 q = P->mods[0].q;
    for (int j = 0; j < N; j++) u0[j] = addmod(u0[j], c0[j], q);
    orc_mod small; mod_init(&small, q, logNs, P->sw.primroot_start);
    u64 *pp[2] = {u0, u1};
    for (int p = 0; p < 2; p++) {
        ntt_inv(&P->mods[0], pp[p]);
        u64 *o = out + (size_t)p * Ns;
        for (int w = 0; w < Ns; w++) o[w] = pp[p][(size_t)w * gap];
        ntt_fwd(&small, o);
    }
    mod_free(&small);
    free(c0); free(c1); free(u0); free(u1);
}

/* ------------------------------------------------------------------------------------------ */
/* path level: fhe.NTT / fhe.Encode                                                           */
/* ------------------------------------------------------------------------------------------ */
typedef struct { const orc_params *P; u64 *data; size_t stride; /* words per element */ int nl; const int *limb_of_lane; } ctuser;
static u64 g_adds, g_subs, g_muls;

static void ct_bfly(nttctx *c, int a, int b) {
    ctuser *u = c->user; const orc_params *P = u->P;
    u64 *x = u->data + (size_t)a * u->stride, *y = u->data + (size_t)b * u->stride;
    if (u->limb_of_lane) {
        for (int l = 0; l < u->nl; l++) {
            u64 q = P->mods[u->limb_of_lane[l]].q, s = x[l], d = y[l];
            x[l] = addmod(s, d, q); y[l] = submod(s, d, q);
        }
        return;
    }
    int N = P->N, k = P->k;
    for (int p = 0; p < 2; p++) for (int i = 0; i < k; i++) {
        u64 q = P->mods[i].q; u64 *xx = x + ((size_t)p * k + i) * N, *yy = y + ((size_t)p * k + i) * N;
        for (int j = 0; j < N; j++) { u64 s = xx[j], d = yy[j]; xx[j] = addmod(s, d, q); yy[j] = submod(s, d, q); }
    }
}
static void ct_mulw(nttctx *c, int a, u64 w) {
    ctuser *u = c->user; const orc_params *P = u->P;
    u64 *x = u->data + (size_t)a * u->stride;
    if (u->limb_of_lane) {
        u64 wi[MAXK], ws[MAXK];
        for (int i = 0; i < P->k; i++) { wi[i] = scalar_mod_qi(P, w, P->mods[i].q); ws[i] = shoup(wi[i], P->mods[i].q); }
        for (int l = 0; l < u->nl; l++) {
            int i = u->limb_of_lane[l];
            x[l] = mulshoup(x[l], wi[i], ws[i], P->mods[i].q);
        }
        return;
    }
    orc_ct_mul_scalar(P, P->k - 1, x, w, x);
}

static void fhe_ntt_generic(ctuser *u, int n, int size, const u64 *roots, int Nfield) {
    int *ids = malloc(sizeof(int) * n);
    for (int i = 0; i < n; i++) ids[i] = i;
    nttctx c = {ct_bfly, ct_mulw, roots, Nfield, u->P->t, 0, 0, 0, u};
    ntt_inner(&c, ids, n, size);
    g_adds = c.n_add; g_subs = c.n_sub; g_muls = c.n_mul;
    /* apply the pointer permutation: result[i] = element ids[i] (cycle-following, one spare) */
    u64 *tmp = malloc(sizeof(u64) * u->stride);
    char *done = calloc(n, 1);
    for (int s = 0; s < n; s++) {
        if (done[s] || ids[s] == s) { done[s] = 1; continue; }
        memcpy(tmp, u->data + (size_t)s * u->stride, sizeof(u64) * u->stride);
        int i = s;
        while (ids[i] != s) {
            memcpy(u->data + (size_t)i * u->stride, u->data + (size_t)ids[i] * u->stride, sizeof(u64) * u->stride);
            done[i] = 1; i = ids[i];
        }
        memcpy(u->data + (size_t)i * u->stride, tmp, sizeof(u64) * u->stride);
        done[i] = 1;
    }
    free(tmp); free(done); free(ids);
}

/* ref: fhe/ntt.go:12-17 */
void orc_fhe_ntt(const orc_params *P, u64 *cts, int n, int size, const u64 *roots, int Nfield) {
    ctuser u = {P, cts, 2 * (size_t)P->k * P->N, 0, NULL};
    fhe_ntt_generic(&u, n, size, roots, Nfield);
}
void orc_fhe_ntt_counts(u64 *a, u64 *s, u64 *m) { *a = g_adds; *s = g_subs; *m = g_muls; }

/* ref: fhe/code.go:8-34 (zeroCol is produced by the host: it is a random encryption) */
void orc_fhe_encode(const orc_params *P, const u64 *matrix, int cols, int rhoInv, const u64 *zero_ct,
                    const u64 *roots, u64 *encoded) {
    size_t ctw = 2 * (size_t)P->k * P->N; int n = cols * rhoInv;
    memcpy(encoded, matrix, sizeof(u64) * ctw * cols);
    for (int i = cols; i < n; i++) memcpy(encoded + ctw * i, zero_ct, sizeof(u64) * ctw);
    orc_fhe_ntt(P, encoded, n, n, roots, n);
}
void orc_fhe_encode_lanes(const orc_params *P, const u64 *matrix_lanes, int cols, int rhoInv,
                          const u64 *zero_lanes, const int *limb_of_lane, int nl, const u64 *roots,
                          u64 *encoded_lanes) {
    int n = cols * rhoInv;
    memcpy(encoded_lanes, matrix_lanes, sizeof(u64) * (size_t)nl * cols);
    for (int i = cols; i < n; i++) memcpy(encoded_lanes + (size_t)nl * i, zero_lanes, sizeof(u64) * nl);
    ctuser u = {P, encoded_lanes, (size_t)nl, nl, limb_of_lane};
    fhe_ntt_generic(&u, n, n, roots, n);
}

/* ------------------------------------------------------------------------------------------ */
/* client side (test infrastructure): PRNG, keys, encoder, encryptor, decryptor               */
/* ------------------------------------------------------------------------------------------ */
typedef struct { u64 s[4]; } rng_t;
static u64 splitmix(u64 *x) { u64 z = (*x += 0x9e3779b97f4a7c15ULL); z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL; return z ^ (z >> 31); }
static void rng_seed(rng_t *r, u64 seed) { for (int i = 0; i < 4; i++) r->s[i] = splitmix(&seed); }
static inline u64 rotl64(u64 x, int k) { return (x << k) | (x >> (64 - k)); }
static u64 rng_next(rng_t *r) {
    u64 *s = r->s, res = rotl64(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl64(s[3], 45);
    return res;
}
static u64 rng_below(rng_t *r, u64 m) { /* unbiased */
    u64 lim = ~0ULL - (~0ULL % m) - 1, x;
    do { x = rng_next(r); } while (x > lim);
    return x % m;
}
static int rng_gauss(rng_t *r) { /* sigma 3.2, bound 19 (lattigo DefaultXe) */
    for (;;) {
        double u1 = ((double)(rng_next(r) >> 11) + 1.0) / 9007199254740993.0, u2 = (double)(rng_next(r) >> 11) / 9007199254740992.0;
        double g = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2) * 3.2;
        long v = lround(g);
        if (v >= -19 && v <= 19) return (int)v;
    }
}
static inline u64 small_mod(int v, u64 q) { return v >= 0 ? (u64)v : q - (u64)(-v); }

void orc_keygen_sk(const orc_params *P, u64 seed, int8_t *sk) {
    /* ternary, Hamming weight min(192, N/2) (lattigo DefaultXs = Ternary{H:192}) */
    rng_t r; rng_seed(&r, seed ^ 0x5ec7e7ULL);
    int N = P->N, h = N / 2 < 192 ? N / 2 : 192;
    memset(sk, 0, N);
    for (int c = 0; c < h;) {
        int pos = (int)rng_below(&r, N);
        if (sk[pos]) continue;
        sk[pos] = (rng_next(&r) & 1) ? 1 : -1; c++;
    }
}
static void small_to_ntt(const orc_params *P, int mi, const int *v, u64 *out) {
    u64 q = P->mods[mi].q;
    for (int j = 0; j < P->N; j++) out[j] = small_mod(v[j], q);
    ntt_fwd(&P->mods[mi], out);
}
static void sk_to_ntt(const orc_params *P, int mi, const int8_t *sk, u64 *out) {
    u64 q = P->mods[mi].q;
    for (int j = 0; j < P->N; j++) out[j] = small_mod(sk[j], q);
    ntt_fwd(&P->mods[mi], out);
}

void orc_keygen_pk(const orc_params *P, const int8_t *sk, u64 seed, u64 *pk) {
    rng_t r; rng_seed(&r, seed ^ 0x9b1cULL);
    int N = P->N, k = P->k;
    int *e = malloc(sizeof(int) * N); u64 *s = malloc(sizeof(u64) * N), *en = malloc(sizeof(u64) * N);
    for (int j = 0; j < N; j++) e[j] = rng_gauss(&r);
    for (int i = 0; i < k; i++) {
        u64 q = P->mods[i].q; u64 *b = pk + (size_t)i * N, *a = pk + ((size_t)k + i) * N;
        for (int j = 0; j < N; j++) a[j] = rng_below(&r, q);
        sk_to_ntt(P, i, sk, s); small_to_ntt(P, i, e, en);
        for (int j = 0; j < N; j++) b[j] = submod(en[j], mulmod(a[j], s[j], q), q); /* -a*s + e */
    }
    free(e); free(s); free(en);
}

/* generic evk: encrypts P * g_d * sIn under sOut (both N ternary coeff vectors given as int arrays) */
static void gen_evk(const orc_params *P, const int8_t *s_in, const u64 *const *s_out_ntt /* [k+np] NTT */, u64 seed, u64 *evk) {
    rng_t r; rng_seed(&r, seed ^ 0xe7aULL);
    int N = P->N, k = P->k, np = P->np, KP = k + np, beta = orc_beta(P);
    int *e = malloc(sizeof(int) * N); u64 *sin = malloc(sizeof(u64) * N), *en = malloc(sizeof(u64) * N);
    for (int d = 0; d < beta; d++) {
        for (int j = 0; j < N; j++) e[j] = rng_gauss(&r);
        for (int m = 0; m < KP; m++) {
            u64 q = P->mods[m].q; u64 R = (u64)((((u128)1) << 64) % q);
            u64 *b = evk + (((size_t)d * 2 + 0) * KP + m) * N, *a = evk + (((size_t)d * 2 + 1) * KP + m) * N;
            for (int j = 0; j < N; j++) a[j] = rng_below(&r, q);
            small_to_ntt(P, m, e, en);
            for (int j = 0; j < N; j++) b[j] = submod(en[j], mulmod(a[j], s_out_ntt[m][j], q), q);
            if (m < k && m >= d * np && m < (d + 1) * np) { /* + P * s_in on the digit's own limbs */
                u64 Pm = 1; for (int x = 0; x < np; x++) Pm = mulmod(Pm, P->mods[k + x].q % q, q);
                sk_to_ntt(P, m, s_in, sin);
                for (int j = 0; j < N; j++) b[j] = addmod(b[j], mulmod(sin[j], Pm, q), q);
            }
            for (int j = 0; j < N; j++) { b[j] = mulmod(b[j], R, q); a[j] = mulmod(a[j], R, q); } /* MForm */
        }
    }
    free(e); free(sin); free(en);
}

/* lattigo: rlwe.KeyGenerator.GenGaloisKey: skOut = pi_{galEl^-1}(sk) */
void orc_keygen_galois(const orc_params *P, const int8_t *sk, u64 galEl, u64 seed, u64 *evk) {
    int N = P->N, KP = P->k + P->np; u64 mask = 2 * (u64)N - 1;
    /* inverse of galEl modulo 2N: the unit group of Z/2N has exponent N/2, and N-1 = -1 mod N/2 */
    u64 inv = 1;
    { u64 e = (u64)N - 1, base = galEl & mask; while (e) { if (e & 1) inv = (inv * base) & mask; base = (base * base) & mask; e >>= 1; } }
    uint32_t *idx = malloc(sizeof(uint32_t) * N);
    orc_automorphism_index(P, inv, idx);
    u64 **so = calloc(KP, sizeof(u64 *)); u64 *tmp = malloc(sizeof(u64) * N);
    for (int m = 0; m < KP; m++) {
        so[m] = malloc(sizeof(u64) * N);
        sk_to_ntt(P, m, sk, tmp);
        for (int j = 0; j < N; j++) so[m][j] = tmp[idx[j]];
    }
    gen_evk(P, sk, (const u64 *const *)so, seed ^ (galEl * 0x100000001b3ULL), evk);
    for (int m = 0; m < KP; m++) free(so[m]);
    free(so); free(tmp); free(idx);
}

/* lattigo: GenEvaluationKey(sk, skSmall): skOut = embedding Y = X^(N/Ns) of the small key */
void orc_keygen_ringswitch(const orc_params *P, const int8_t *sk, const int8_t *sk_small, int logNs, u64 seed, u64 *evk) {
    int N = P->N, Ns = 1 << logNs, gap = N / Ns, KP = P->k + P->np;
    int8_t *emb = calloc(N, 1);
    for (int j = 0; j < Ns; j++) emb[(size_t)j * gap] = sk_small[j];
    u64 **so = calloc(KP, sizeof(u64 *));
    for (int m = 0; m < KP; m++) { so[m] = malloc(sizeof(u64) * N); sk_to_ntt(P, m, emb, so[m]); }
    gen_evk(P, sk, (const u64 *const *)so, seed ^ 0x51ULL, evk);
    for (int m = 0; m < KP; m++) free(so[m]);
    free(so); free(emb);
}

/* lattigo: bgv.Encoder.Encode (EncodeRingT + RingT2Q(scaleUp) + NTT), scale = 1 */
void orc_encode_pt(const orc_params *P, const u64 *values, int n, int level, u64 *pt) {
    int N = P->N; u64 t = P->t;
    u64 *pT = calloc(N, sizeof(u64));
    for (int i = 0; i < n && i < N; i++) pT[P->perm[i]] = values[i] % t;
    ntt_inv(&P->modt, pT);
    for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q, tinv = invmod(t % q, q); u64 *o = pt + (size_t)i * N;
        for (int j = 0; j < N; j++) o[j] = mulmod(pT[j] % q, tinv, q);
        ntt_fwd(&P->mods[i], o);
    }
    free(pT);
}

/* lattigo: rlwe.Encryptor with pk (restricted to Q): ct = (u*pk0 + e0 + pt, u*pk1 + e1) */
void orc_encrypt_pk(const orc_params *P, const u64 *pk, const u64 *pt, u64 seed, u64 *ct) {
    rng_t r; rng_seed(&r, seed ^ 0xc1f3ULL);
    int N = P->N, k = P->k;
    int *u = malloc(sizeof(int) * N), *e0 = malloc(sizeof(int) * N), *e1 = malloc(sizeof(int) * N);
    u64 *un = malloc(sizeof(u64) * N), *en = malloc(sizeof(u64) * N);
    for (int j = 0; j < N; j++) { u64 x = rng_next(&r) % 3; u[j] = x == 0 ? 0 : (x == 1 ? 1 : -1); }
    for (int j = 0; j < N; j++) { e0[j] = rng_gauss(&r); e1[j] = rng_gauss(&r); }
    for (int i = 0; i < k; i++) {
        u64 q = P->mods[i].q; u64 *c0 = ct + (size_t)i * N, *c1 = ct + ((size_t)k + i) * N;
        small_to_ntt(P, i, u, un);
        small_to_ntt(P, i, e0, en);
        for (int j = 0; j < N; j++) {
            u64 v = addmod(mulmod(un[j], pk[(size_t)i * N + j], q), en[j], q);
            c0[j] = pt ? addmod(v, pt[(size_t)i * N + j], q) : v;
        }
        small_to_ntt(P, i, e1, en);
        for (int j = 0; j < N; j++) c1[j] = addmod(mulmod(un[j], pk[((size_t)k + i) * N + j], q), en[j], q);
    }
    free(u); free(e0); free(e1); free(un); free(en);
}

/* Garner mixed-radix: residues x[0..L] mod q_0..q_L -> X mod m for X in [0,Q) */
static void crt_to_mod(const orc_params *P, int level, const u64 *const *limbs, int N, u64 m, int add_half, u64 *out) {
    int L = level + 1;
    u64 inv[MAXK][MAXK]; /* inv[i][j] = q_j^-1 mod q_i, j<i */
    for (int i = 0; i < L; i++) for (int j = 0; j < i; j++) inv[i][j] = invmod(P->mods[j].q % P->mods[i].q, P->mods[i].q);
    /* floor(Q/2) = (Q-1)/2: residue mod q_i is (q_i-1)/2; mod m is (Q mod m - 1)/2 */
    u64 Qm = 1 % m; for (int i = 0; i < L; i++) Qm = mulmod(Qm, P->mods[i].q % m, m);
    u64 half_m = mulmod(submod(Qm, 1 % m, m), invmod(2, m), m);
    for (int c = 0; c < N; c++) {
        u64 v[MAXK];
        for (int i = 0; i < L; i++) {
            u64 qi = P->mods[i].q, x = limbs[i][c];
            if (add_half) x = addmod(x, (qi - 1) >> 1, qi);
            /* v_i = (x - (v_0 + v_1 q_0 + ...)) / (q_0..q_{i-1}) mod q_i */
            u64 tcur = x;
            for (int j = 0; j < i; j++) tcur = mulmod(submod(tcur, v[j] % qi, qi), inv[i][j], qi);
            v[i] = tcur;
        }
        u64 acc = 0;
        for (int i = L - 1; i >= 0; i--) acc = addmod(mulmod(acc, P->mods[i].q % m, m), v[i] % m, m);
        out[c] = add_half ? submod(acc, half_m, m) : acc;
    }
}

/* lattigo: rlwe.Decryptor.DecryptNew + bgv.Encoder.Decode (RingQ2T scaleDown, DecodeRingT) */
void orc_decrypt_decode(const orc_params *P, const int8_t *sk, const u64 *ct, int level, u64 scale, u64 *values) {
    int N = P->N; u64 t = P->t;
    u64 *buf = malloc(sizeof(u64) * (size_t)(level + 1) * N), *s = malloc(sizeof(u64) * N);
    const u64 *limbs[MAXK];
    for (int i = 0; i <= level; i++) {
        u64 q = P->mods[i].q; u64 *o = buf + (size_t)i * N;
        const u64 *c0 = LIMB(ct, level, N, 0, i), *c1 = LIMB(ct, level, N, 1, i);
        sk_to_ntt(P, i, sk, s);
        for (int j = 0; j < N; j++) o[j] = addmod(c0[j], mulmod(c1[j], s[j], q), q);
        ntt_inv(&P->mods[i], o);
        u64 tq = t % q;
        for (int j = 0; j < N; j++) o[j] = mulmod(o[j], tq, q);
        limbs[i] = o;
    }
    u64 *pT = malloc(sizeof(u64) * N);
    crt_to_mod(P, level, limbs, N, t, 1, pT);
    u64 sinv = invmod(scale % t, t);
    for (int j = 0; j < N; j++) pT[j] = mulmod(pT[j], sinv, t);
    ntt_fwd(&P->modt, pT);
    for (int i = 0; i < N; i++) values[i] = pT[P->perm[i]];
    free(buf); free(s); free(pT);
}

/* diagnostic: log2 of the infinity norm of the centred value (c0 + c1 s) mod Q_level  (noise + T^-1 m term);
   for an encryption of zero this is the noise size in bits.  Uses long double on the mixed-radix digits. */
double orc_noise_bits(const orc_params *P, const int8_t *sk, const u64 *ct, int level) {
    int N = P->N, L = level + 1;
    u64 *buf = malloc(sizeof(u64) * (size_t)L * N), *s = malloc(sizeof(u64) * N);
    for (int i = 0; i < L; i++) {
        u64 q = P->mods[i].q; u64 *o = buf + (size_t)i * N;
        const u64 *c0 = LIMB(ct, level, N, 0, i), *c1 = LIMB(ct, level, N, 1, i);
        sk_to_ntt(P, i, sk, s);
        for (int j = 0; j < N; j++) o[j] = addmod(c0[j], mulmod(c1[j], s[j], q), q);
        ntt_inv(&P->mods[i], o);
    }
    u64 inv[MAXK][MAXK];
    for (int i = 0; i < L; i++) for (int j = 0; j < i; j++) inv[i][j] = invmod(P->mods[j].q % P->mods[i].q, P->mods[i].q);
    double best = -1e9;
    for (int c = 0; c < N; c++) {
        u64 v[MAXK];
        for (int i = 0; i < L; i++) {
            u64 qi = P->mods[i].q, x = addmod(buf[(size_t)i * N + c], (qi - 1) >> 1, qi), tcur = x;
            for (int j = 0; j < i; j++) tcur = mulmod(submod(tcur, v[j] % qi, qi), inv[i][j], qi);
            v[i] = tcur;
        }
        /* centred value = sum_i (v_i - h_i) * prod_{j<i} q_j with h_i = (q_i-1)/2 (telescoping digits of (Q-1)/2):
           magnitude ~ |v_top - h_top| * prod_{j<top} q_j for the highest differing digit */
        double b = 0.0;
        for (int i = L - 1; i >= 0; i--) {
            u64 h = (P->mods[i].q - 1) >> 1;
            if (v[i] != h) {
                double d = v[i] > h ? (double)(v[i] - h) : (double)(h - v[i]);
                b = log2(d);
                for (int j = 0; j < i; j++) b += log2((double)P->mods[j].q);
                break;
            }
        }
        if (b > best) best = b;
    }
    free(buf); free(s);
    return best;
}
