// tests/emul/emul_matcher.cpp -- TEST INFRASTRUCTURE (never shipped, never a product path).
//
// Compiles the DEVICE matcher header (baitstr_b200/csrc/matcher.cuh) for the host with
// g++ by shimming the handful of CUDA integer intrinsics it uses, so the exact device
// source can be fuzzed against oracle/str_oracle.c in the GPU-less build container before
// GPU time is spent.  The product library has no CPU path; this object is only ever
// loaded by tests/test_matcher_emul.py.
#include <stdint.h>
#include <string.h>
#include <stdlib.h>

#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ inline

static inline uint32_t __funnelshift_r(uint32_t lo, uint32_t hi, uint32_t s)
{
    const uint64_t v = ((uint64_t)hi << 32) | lo;
    return (uint32_t)(v >> (s & 31));
}
static inline uint32_t __funnelshift_l(uint32_t lo, uint32_t hi, uint32_t s)
{
    const uint64_t v = ((uint64_t)hi << 32) | lo;
    return (uint32_t)((v << (s & 31)) >> 32);
}
static inline uint32_t __brev(uint32_t x)
{
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
    return __builtin_bswap32(x);
}
static inline int __clz(uint32_t x) { return x ? __builtin_clz(x) : 32; }
static inline int __ffs(uint32_t x) { return __builtin_ffs((int)x); }
static inline int __popc(uint32_t x) { return __builtin_popcount(x); }

#include "../../baitstr_b200/csrc/matcher.cuh"

extern "C" {

struct emul_params { int n, f, kmask, rmhm, altalgo; };

// results: 4 x uint32 per read {fstart, rstart, fck, rck}; same convention as k_match_reads
// (forward hit reported even when the reverse strand has none).
// force_w: 0 = dispatch like the device (window_words), else force the width 2/4/6/8 or 9 (bytes path);
// 6 goes through the WMAX = 6 instantiation (the 6-word window exists only there)
void emul_match_reads(const uint8_t *seqs, const uint64_t *off, uint64_t n, const emul_params *ep,
                      int force_w, uint32_t *res)
{
    bstr::ScanParams prm;
    prm.n = ep->n < 1 ? 1 : ep->n; prm.f = ep->f < 0 ? 0 : ep->f; prm.kmask = ep->kmask;
    prm.rmhm = ep->rmhm; prm.altalgo = ep->altalgo;
    for (uint64_t r = 0; r < n; r++) {
        const int L = (int)(off[r + 1] - off[r]);
        // private copy with the slack the device buffers guarantee (64 bytes each side)
        uint8_t *buf = (uint8_t *)calloc((size_t)L + 192, 1);
        uint8_t *s = buf + 64 + (off[r] & 3);              // keep the original misalignment
        memcpy(s, seqs + off[r], (size_t)L);
        bstr::Hit f, rc;
        int w = bstr::window_words(L, prm);
        if (force_w && force_w >= w) w = force_w;
        if (force_w == 6) bstr::match_read<6>(s, L, prm, w, f, rc);
        else bstr::match_read(s, L, prm, w, f, rc);
        const bool sel = f.k && rc.k;
        res[4 * r + 0] = f.k ? (uint32_t)f.start : 0;
        res[4 * r + 1] = sel ? (uint32_t)rc.start : 0;
        res[4 * r + 2] = f.k ? ((uint32_t)f.copies << 3 | (uint32_t)f.k) : 0;
        res[4 * r + 3] = sel ? ((uint32_t)rc.copies << 3 | (uint32_t)rc.k) : 0;
        free(buf);
    }
}

// -d report of one read through the device code (matcher.cuh: debug_read); codes_fwd / codes_rc
// need bstr::debug_codes_cap bytes each
void emul_debug_read(const uint8_t *seq, int L, const emul_params *ep, uint8_t *codes_fwd, uint8_t *codes_rc,
                     uint32_t *info)
{
    bstr::ScanParams prm;
    prm.n = ep->n < 1 ? 1 : ep->n; prm.f = ep->f < 0 ? 0 : ep->f; prm.kmask = ep->kmask;
    prm.rmhm = ep->rmhm; prm.altalgo = ep->altalgo;
    uint8_t *buf = (uint8_t *)calloc((size_t)L + 192, 1);
    memcpy(buf + 64, seq, (size_t)L);
    uint32_t o[8];
    bstr::debug_read(buf + 64, L, prm, codes_fwd, codes_rc, o);
    for (int i = 0; i < 8; i++) info[i] = o[i];
    free(buf);
}
uint32_t emul_debug_codes_cap(int L, int n, int kmask) { return bstr::debug_codes_cap(L, n, kmask); }

// exhaustive check hook for the SWAR base validator: returns 1 if the device code calls the
// L bytes at p (any alignment) invalid
int emul_bases_invalid(const uint8_t *p, int L)
{
    uint8_t *buf = (uint8_t *)calloc((size_t)L + 64, 1);
    memset(buf, 0xEE, (size_t)L + 64);                 // poison around the range
    uint8_t *s = buf + 16 + ((uintptr_t)p & 3);
    memcpy(s, p, (size_t)L);
    const int r = bstr::bases_invalid(s, L) ? 1 : 0;
    free(buf);
    return r;
}
}
