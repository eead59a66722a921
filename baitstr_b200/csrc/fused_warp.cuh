// fused_warp.cuh -- the single-pass kernel of the select stage (sm_100a), warp-autonomous.
//
// Every WARP owns a small tile of raw FASTQ (about 30 records plus a look-ahead for the
// record that straddles the tile end) in its private slice of shared memory and takes it
// through the whole stage without ever meeting another warp at a CTA barrier:
//   fastq.next()            /root/reference/src/fastq.pyx:72-94            (line split, checks)
//   ProcessSequence         /root/reference/src/select_STR_reads.pyx:294-319 (both strands)
//   PrintSequence           /root/reference/src/select_STR_reads.pyx:171-185 (annotated FASTQ)
//   ordered output          /root/reference/src/select_STR_reads.pyx:363-367 (input order)
// Warps are persistent: each pulls the next tile ticket, so a tile only ever waits (in the
// two decoupled look-backs: line count, output bytes) for tiles that already run.  The
// CTA-wide version of this kernel (fused.cuh) spent 45 % of its warp samples at
// __syncthreads (profiles/r01_ncu_fused_v3_phases.txt): one warp doing a look-back or a
// long match kept the other seven waiting.  Here the only synchronisation is __syncwarp.
// HBM traffic stays the algorithmic minimum: every input byte read once (cp.async, 16 B),
// every output byte written once (16 B stores from the staging slice).
#pragma once
#include <stdint.h>
#include "fused.cuh"

namespace bstr {

constexpr int FW_WARPS = 4;                    // warps per CTA (independent of each other)
constexpr int FW_THREADS = 32 * FW_WARPS;
constexpr int FW_MAXCPT = 32;                  // 16-byte chunks per lane in the index phase
constexpr int FW_MAXREGION = 512 * FW_MAXCPT;  // 16 KiB staged per warp
constexpr int FW_RECMAX = 64;                  // records a tile may own (two rounds of 32)
constexpr int FW_MINTILE = 2048;
constexpr int FW_MAXSMEM = 227 * 1024;

struct WarpGeom {
    uint32_t tile;           // bytes owned by one warp tile (multiple of 16)
    uint32_t region;         // tile + look-ahead (multiple of 16, <= FW_MAXREGION)
    uint32_t chunks, cpt;    // region / 16; chunks per lane (contiguous)
    uint32_t nlmax, stage;
    uint32_t off_in, off_stage, off_rec, off_outoff, off_nl;   // byte offsets in the warp's slice
    uint32_t warp_bytes;
};

inline WarpGeom make_warp_geometry(uint32_t tile, uint32_t ovl, uint32_t stage)
{
    WarpGeom g;
    g.tile = tile; g.region = tile + ovl; g.chunks = g.region / 16; g.cpt = (g.chunks + 31) / 32;
    g.nlmax = 5 * FW_RECMAX + 64; g.stage = stage;
    uint32_t o = 0;
    g.off_in = o;       o += 16 + g.region + 32;
    g.off_stage = o;    o += g.stage + 32;
    g.off_rec = o;      o += FW_RECMAX * 16;
    g.off_outoff = o;   o += FW_RECMAX * 4;
    g.off_nl = o;       o += (g.nlmax + 8) * 2;
    g.warp_bytes = (o + 15) & ~15u;
    return g;
}

__device__ __forceinline__ void cp_async16_zfill(uint8_t *smem_dst, const uint8_t *gsrc, uint32_t src_bytes)
{
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// Line index of this lane's chunks (see index_chunks in fused.cuh), FW_MAXCPT chunks per lane.
template <bool CAREFUL>
__device__ __forceinline__ uint32_t warp_index_chunks(const uint8_t *sin, const WarpGeom &g, uint32_t c0,
                                                      uint64_t ts, uint64_t n, int final,
                                                      uint64_t (&mk)[FW_MAXCPT / 4])
{
    uint32_t cnt = 0;
    const uint32_t own = g.tile / 16;
#pragma unroll
    for (int i = 0; i < FW_MAXCPT; i++) {
        if (i < (int)g.cpt) {
            const uint32_t c = c0 + i;
            uint32_t m = 0;
            if (c < g.chunks) {
                m = nl_mask_of(*reinterpret_cast<const uint4 *>(sin + 16 * c));
                if (CAREFUL) {
                    const uint64_t gpos = ts + 16ull * c;
                    if (gpos >= n) m = 0;
                    else {
                        const uint64_t rem = n - gpos;
                        if (rem < 16) m &= (1u << rem) - 1u;
                        // an unterminated last line: its last byte is dropped by line[:-1]
                        // (fastq.pyx:78,83,90), i.e. it plays the newline
                        if (final && rem <= 16) m |= 1u << (rem - 1);
                    }
                }
            }
            mk[i >> 2] |= (uint64_t)m << (16 * (i & 3));
            const uint32_t pc = __popc(m);
            cnt += (pc << 16) | (c < own ? pc : 0u);
        }
    }
    return cnt;
}

__global__ void __launch_bounds__(FW_THREADS, 3)
k_fused_warp(const uint8_t *__restrict__ in, uint64_t n, int final, ScanParams prm, int illumina,
             BatchState *st, unsigned long long *lb_lines, unsigned long long *lb_out,
             unsigned long long *ticket, uint32_t ntiles, ReadResult *res, uint8_t *out,
             uint64_t outcap, int exp_mask, const WarpGeom g)
{
    extern __shared__ __align__(16) uint8_t fused_smem[];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint8_t *const slice = fused_smem + (size_t)wid * g.warp_bytes;
    uint8_t *const in_raw = slice + g.off_in;
    uint8_t *const sin = in_raw + 16;                        // tile bytes; 16 B of zero slack before
    uint8_t *const stage = slice + g.off_stage;
    ReadResult *const s_rec = reinterpret_cast<ReadResult *>(slice + g.off_rec);
    uint32_t *const s_outoff = reinterpret_cast<uint32_t *>(slice + g.off_outoff);
    uint16_t *const s_nl = reinterpret_cast<uint16_t *>(slice + g.off_nl);   // newline position + 1
    if (lane < 4) {
        reinterpret_cast<uint32_t *>(in_raw)[lane] = 0u;
        reinterpret_cast<uint32_t *>(sin + g.region)[lane] = 0u;
        reinterpret_cast<uint32_t *>(sin + g.region + 16)[lane] = 0u;
    }

    for (;;) {
        __syncwarp();                                    // the previous tile's smem reads are done
        uint32_t tile = 0;
        if (lane == 0) tile = (uint32_t)atomicAdd(ticket, 1ull);
        tile = __shfl_sync(0xFFFFFFFFu, tile, 0);
        if (tile >= ntiles) break;
        const uint64_t ts = (uint64_t)tile * g.tile;
        const bool reach = ts + g.region >= n;           // the staged window covers the batch end
        const bool last = tile == ntiles - 1;

        // ---- A: stage the tile (+ look-ahead) in shared memory: cp.async, 16 B, zero fill ---
        for (uint32_t c = lane; c < g.chunks; c += 32) {
            const uint64_t off = ts + 16ull * c;
            const uint32_t nb = off < n ? (n - off < 16 ? (uint32_t)(n - off) : 16u) : 0u;
            cp_async16_zfill(sin + 16 * c, nb ? in + off : in, nb);
        }
        cp_async_wait_all();
        __syncwarp();

        // ---- B: line index of the window -------------------------------------------------
        uint64_t mk[FW_MAXCPT / 4] = {0, 0, 0, 0, 0, 0, 0, 0};
        const uint32_t c0 = lane * g.cpt;
        const uint32_t cnt = reach ? warp_index_chunks<true>(sin, g, c0, ts, n, final, mk)
                                   : warp_index_chunks<false>(sin, g, c0, ts, n, final, mk);   // all << 16 | own
        uint32_t inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if (lane >= d) inc += t;
        }
        const uint32_t tot = __shfl_sync(0xFFFFFFFFu, inc, 31);
        const uint32_t ex = inc - cnt;
        const uint32_t shift = tile == 0 ? 1u : 0u;      // tile 0: virtual newline before byte 0
        const uint32_t real_own = tot & 0xFFFFu;
        uint32_t cnt_own = real_own + shift, cnt_all = (tot >> 16) + shift;
        bool bail = cnt_all + 2 > g.nlmax;
        {
            uint32_t idx = (ex >> 16) + shift;
#pragma unroll
            for (int w = 0; w < FW_MAXCPT / 4; w++) {
                uint64_t m = mk[w];
                const uint32_t base = 16u * (c0 + 4 * w) + 1u;
                while (m) {
                    const int b = __ffsll((long long)m) - 1;
                    m &= m - 1;
                    if (idx < g.nlmax) s_nl[idx] = (uint16_t)(base + b);
                    idx++;
                }
            }
            if (lane == 0 && shift) s_nl[0] = 0;
        }

        // ---- C: global line number of the tile's first line (look-back #1) -----------------
        const unsigned long long lbase = warp_lookback(lb_lines, (int)tile, real_own);
        const long long gbase = tile == 0 ? -1ll : (long long)lbase;
        const long long G = gbase + (long long)cnt_all;  // lines of the whole batch, if `reach`
        if (final && reach && (G & 3) == 3 && !bail) {
            // EOF after the separator line: accepted with an empty quality string
            // (readline() returns '' at fastq.pyx:89)
            if (lane == 0) s_nl[cnt_all] = (uint16_t)(n - ts + 1);
            cnt_all++;
        }
        __syncwarp();                                    // s_nl complete (virtual entry included)
        if (last && lane == 0) {
            const unsigned long long lines = (unsigned long long)G + ((final && (G & 3) == 3) ? 1ull : 0ull);
            st->n_lines = lines;
            st->n_records = lines >> 2;
            if (final) {
                if ((G & 3) == 1 || (G & 3) == 2)        // file ends inside a record (fastq.pyx:86)
                    atomicMin(&st->first_bad, ((unsigned long long)(G >> 2) << 8) | BAD_TRUNCATED);
                atomicMax(&st->consumed, (unsigned long long)n);
            }
        }
        const uint32_t i0 = (uint32_t)((3 - gbase) & 3);
        const uint32_t nrec_own = cnt_own > i0 ? (cnt_own - i0 + 3) / 4 : 0u;
        uint32_t nproc = cnt_all >= i0 + 5 ? (cnt_all - 1 - i0) / 4 : 0u;
        if (nproc > nrec_own) nproc = nrec_own;
        if (nproc < nrec_own && !reach) bail = true;      // a record runs past the look-ahead
        if (nproc > FW_RECMAX) bail = true;
        if (!bail && nrec_own && lane == 0) {
            const uint32_t iL = i0 + 4 * (nrec_own - 1);
            if (!(shift && iL == 0)) atomicMax(&st->consumed, ts + s_nl[iL]);
        }
        if (!bail && final && reach && nproc < nrec_own && lane == 0) {
            // a record cut short by EOF: the reference still tests its header (fastq.pyx:77) and a
            // complete second line (fastq.pyx:81) before the missing line raises (fastq.pyx:86)
            const uint32_t i = i0 + 4 * nproc, after = cnt_all - 1 - i;
            if (after >= 1) {
                const unsigned long long r = (unsigned long long)(G >> 2);
                if (sin[s_nl[i]] != '@') atomicMin(&st->first_bad, (r << 8) | BAD_HEADER);
                else if (after >= 2 && bases_invalid(sin + s_nl[i + 1], (int)s_nl[i + 2] - 1 - (int)s_nl[i + 1]))
                    atomicMin(&st->first_bad, (r << 8) | BAD_BASES);
            }
        }
        if (bail) nproc = 0;

        // ---- D: per record: fastq.next() checks + both-strand match (lane per record) ---------
        uint32_t nsel = 0;
#pragma unroll 1
        for (uint32_t jb = 0; jb < FW_RECMAX; jb += 32) {
            const uint32_t j = jb + lane;
            if (jb >= nproc) {                            // warp-uniform: nothing left in this round
                s_outoff[j] = 0;
                continue;
            }
            const bool live = j < nproc;
            RecordLines ln = {};
            int L = 0;
            if (live) { ln = tile_record_lines(s_nl, (int)(i0 + 4 * j)); L = (int)(ln.e1 - ln.s1); }
            const int wmax = __reduce_max_sync(0xFFFFFFFFu, (unsigned)window_words(L, prm));
            Hit f, rc;
            f.k = rc.k = 0; f.copies = rc.copies = f.start = rc.start = 0;
            bool sel = false;
            uint32_t olen = 0;
            if (live) {
                const unsigned long long r = (unsigned long long)((gbase + (long long)(i0 + 4 * j) + 1) >> 2);
                int bad = 0;
                if (sin[ln.s0] != '@') bad = BAD_HEADER;                       // fastq.pyx:77
                else if (!(exp_mask & 2) && bases_invalid(sin + ln.s1, L)) bad = BAD_BASES;        // fastq.pyx:81
                else if (sin[ln.s2] != '+') bad = BAD_SEPARATOR;                // fastq.pyx:86
                if (bad) atomicMin(&st->first_bad, (r << 8) | (unsigned)bad);
                else if (!(exp_mask & 1)) match_read(sin + ln.s1, L, prm, wmax, f, rc);
                sel = f.k && rc.k;
                ReadResult o;
                o.fstart = sel ? f.start : 0; o.rstart = sel ? rc.start : 0;
                o.fck = sel ? ((uint32_t)f.copies << 3 | f.k) : 0;
                o.rck = sel ? ((uint32_t)rc.copies << 3 | rc.k) : 0;
                s_rec[j] = o;
                if (res) res[r] = o;
                if (sel) olen = record_out_len(ln.e0 - ln.s0 - 1, L, ln.e3 - ln.s3, f, rc);
            }
            s_outoff[j] = olen;
            nsel += __popc(__ballot_sync(0xFFFFFFFFu, sel));
        }
        __syncwarp();

        // ---- E: output offsets: records inside the tile, then the tile (look-back #2) ---------
        const uint32_t o0 = s_outoff[2 * lane], o1 = s_outoff[2 * lane + 1];
        uint32_t oinc = o0 + o1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, oinc, d);
            if (lane >= d) oinc += t;
        }
        uint32_t total = __shfl_sync(0xFFFFFFFFu, oinc, 31);
        const uint32_t oex = oinc - (o0 + o1);
        __syncwarp();
        s_outoff[2 * lane] = oex;
        s_outoff[2 * lane + 1] = oex + o0;
        if (total > g.stage - 16) { bail = true; total = 0; nproc = 0; }
        if (bail && lane == 0) atomicExch(&st->fallback, 1ull);
        const unsigned long long out_base = warp_lookback(lb_out, (int)tile, total);
        if (lane == 0) {
            if (last) st->out_bytes = out_base + total;
            if (nsel && !bail) atomicAdd(&st->n_selected, (unsigned long long)nsel);
        }
        __syncwarp();
        if (total == 0 || out_base + total > outcap) continue;   // host grows the buffer and re-runs

        // ---- F: PrintSequence into the staging slice (same 16-byte phase as the destination) --
        const uint32_t a0 = (uint32_t)(out_base & 15);
        for (uint32_t j = lane; j < nproc && !(exp_mask & 4); j += 32) {
            const ReadResult rr = s_rec[j];
            if (rr.fck == 0) continue;
            const RecordLines ln = tile_record_lines(s_nl, (int)(i0 + 4 * j));
            format_record_words(sin, ln, rr, illumina, stage + a0 + s_outoff[j]);
        }
        __syncwarp();

        // ---- G: flush the staging slice with aligned 16-byte stores ---------------------------
        if (exp_mask & 8) continue;
        uint8_t *dst = out + out_base;
        const uint32_t head = min(total, (16u - a0) & 15u);
        if (lane < head) dst[lane] = stage[a0 + lane];
        const uint32_t nchunk = (total - head) >> 4;
        const uint8_t *sb = stage + a0 + head;
        for (uint32_t c = lane; c < nchunk; c += 32)
            *reinterpret_cast<uint4 *>(dst + head + 16 * c) = *reinterpret_cast<const uint4 *>(sb + 16 * c);
        const uint32_t done = head + (nchunk << 4);
        if (lane < total - done) dst[done + lane] = stage[a0 + done + lane];
    }
}

}  // namespace bstr
