// Host-side 64-bit modular helpers used to build the device constant tables.
#pragma once
#include <cstdint>

namespace lmn {
typedef unsigned long long u64;
typedef unsigned __int128 u128;

inline u64 h_mulmod(u64 a, u64 b, u64 m) { return (u64)(((u128)a * b) % m); }
inline u64 h_addmod(u64 a, u64 b, u64 m) { u64 r = a + b; return (r >= m || r < a) ? r - m : r; }
inline u64 h_submod(u64 a, u64 b, u64 m) { return a >= b ? a - b : a + m - b; }
inline u64 h_powmod(u64 a, u64 e, u64 m) {
    u64 r = 1 % m; a %= m;
    while (e) { if (e & 1) r = h_mulmod(r, a, m); a = h_mulmod(a, a, m); e >>= 1; }
    return r;
}
inline u64 h_invmod(u64 a, u64 m) { return h_powmod(a % m, m - 2, m); }  // m prime
inline u64 h_shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
inline u64 h_mform(u64 a, u64 q) { return (u64)(((u128)a << 64) % q); }   // a * 2^64 mod q
inline u64 h_qinv64(u64 q) { u64 x = 1; for (int i = 0; i < 63; i++) { x *= q; q *= q; } return x; }  // q^-1 mod 2^64
inline u64 h_bitrev(u64 x, int bits) {
    u64 r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}
inline bool h_is_pow2(u64 x) { return x && !(x & (x - 1)); }
inline int h_log2(u64 x) { int l = 0; while ((1ull << l) < x) l++; return l; }
}  // namespace lmn
