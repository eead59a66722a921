// merge.cu -- bstr_merge_reads: the merge stage (merge_STR_reads) on the GPU, first slice
// (SURVEY.md 8(f) rank 2).  Replaces MergeShortTandemRepeatReads, /root/reference/src/merge_STR_reads.c:308-509.
//
// The reference runs one loop over the reads; a read meets earlier reads only through the map keyed
// by "motif lflank rflank", looked up for the read as it is and reverse-complemented.  Reads whose
// keys never touch cannot influence each other, so the reads are cut into CHAINS (connected
// components over the two keys of every read, merge_plan.h) and every chain is walked in input
// order by ONE WARP: the 32 lanes fill the two local-alignment matrices of AlignFlanks along
// anti-diagonals, lane 0 walks the traceback and rewrites the block (merge_step.cuh, merge_core.cuh
// -- the same source the CPU suite runs on the host).  Blocks live in HBM, one slot per key; the
// two-bit traceback matrix, the three live anti-diagonals and the consensus flanks in shared memory.
// Parsing, keys, chains, the reference's table order and printing are host work (merge_plan.h).
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <memory>
#include <numeric>

#include "host_common.h"
#include "merge_plan.h"
#include "merge_step.cuh"

namespace bmerge {

// Shared memory of one warp for a geometry (longest flank maxf, consensus room cap): the two-bit traceback
// matrix, three anti-diagonals of scores, four consensus strings, the two flanks being aligned
struct Geometry {
    int maxf, pitch, cap;
    uint32_t o_diag, o_cons, o_seq, warp_bytes;
    int warps, ctas_per_sm;
};
static Geometry geometry_for(int maxf)
{
    auto up16 = [](uint32_t x) { return (x + 15u) & ~15u; };
    Geometry g;
    g.maxf = maxf < 31 ? 31 : (maxf > MAXF ? MAXF : maxf);
    g.pitch = pitch_for(g.maxf);
    g.cap = 2 * g.maxf + 64 > MAXB ? MAXB : (int)up16(2 * g.maxf + 64);
    g.o_diag = (uint32_t)t_bytes_for(g.maxf);
    g.o_cons = g.o_diag + up16(3u * (g.maxf + 1) * 2u);
    g.o_seq = g.o_cons + 4u * (uint32_t)g.cap;
    g.warp_bytes = g.o_seq + up16(2u * (g.maxf + 1));
    const uint32_t sm = 227u * 1024u;
    g.warps = (int)std::min<uint32_t>(8u, sm / g.warp_bytes);
    g.ctas_per_sm = (int)std::max<uint32_t>(1u, std::min<uint32_t>(8u, sm / (g.warps * g.warp_bytes + 1024u)));
    return g;
}

struct WarpExec {
    int lane, pitch;
    int16_t *d0, *d1, *d2;
    char *s1, *s2;
    // Align's matrix (:86-110) along anti-diagonals d = i + j: a cell needs (i, j-1) and (i-1, j) of
    // diagonal d-1 and (i-1, j-1) of d-2, all indexed by their row here
    __device__ __forceinline__ uint64_t dp_fill(const char *t1, int len1, const char *t2, int len2, uint8_t *T)
    {
        uint64_t best = 0;
        int16_t *p2 = d0, *p1 = d1, *cur = d2;
        if (len1 >= 2 && len2 >= 2) {
            for (int a = lane; a < len1 - 1; a += 32) s1[a] = t1[a];         // the flanks into shared memory: every cell reads both
            for (int a = lane; a < len2 - 1; a += 32) s2[a] = t2[a];
            __syncwarp();
            for (int d = 2; d <= len1 + len2 - 2; d++) {
                const int ilo = max(1, d - (len2 - 1)), ihi = min(len1 - 1, d - 1);
                for (int i = ilo + lane; i <= ihi; i += 32) {
                    const int j = d - i;
                    const int left = j >= 2 ? p1[i] : 0, up = i >= 2 ? p1[i - 1] : 0, dg = (i >= 2 && j >= 2) ? p2[i - 1] : 0;
                    unsigned code;
                    const int a = dp_cell(left, up, dg, s1[i - 1] == s2[j - 1], &code);
                    cur[i] = (int16_t)a;
                    t_put(T, pitch, i, j, code);     // cells of one diagonal lie pitch - 1 apart: never in the same byte, an odd number of words apart
                    const uint64_t key = best_key(a, i, j);
                    best = key > best ? key : best;
                }
                __syncwarp();
                int16_t *t = p2; p2 = p1; p1 = cur; cur = t;
            }
        }
        for (int o = 16; o; o >>= 1) {
            const uint64_t other = __shfl_xor_sync(0xFFFFFFFFu, best, o);
            best = other > best ? other : best;
        }
        return best;
    }
    __device__ __forceinline__ bool leader() const { return lane == 0; }
    __device__ __forceinline__ int bcast(int v) const { return __shfl_sync(0xFFFFFFFFu, v, 0); }
    __device__ __forceinline__ void sync() const { __syncwarp(); }
};

// a warp per chain, chains handed out by a ticket (longest first: the host sorts them).  With a redo list the
// geometry is a small one: a chain that outgrows it (or fails in any way) is put on the list and walked
// again by a second launch with the largest geometry, which first takes the chain's blocks back (`reset`)
__global__ void __launch_bounds__(256)
k_merge_chains(const Read *reads, const uint32_t *chain_start, const uint32_t *chain_reads, const uint32_t *work, uint32_t nwork,
               BlockHead *heads, char *bseqs, char *bquals, const char *fwd, const char *qfwd, const char *rev, const char *qrev,
               unsigned *ticket, int *err, uint32_t *redo_list, unsigned *redo_count, int reset, Geometry g)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t *w = smem + (size_t)(threadIdx.x >> 5) * g.warp_bytes;
    int16_t *diag = reinterpret_cast<int16_t *>(w + g.o_diag);
    char *cons = reinterpret_cast<char *>(w + g.o_cons), *seq = reinterpret_cast<char *>(w + g.o_seq);
    WarpExec x{(int)(threadIdx.x & 31), g.pitch, diag, diag + (g.maxf + 1), diag + 2 * (g.maxf + 1), seq, seq + (g.maxf + 1)};
    const Scratch s{w, cons, cons + g.cap, cons + 2 * g.cap, cons + 3 * g.cap, g.maxf, g.pitch, g.cap};
    for (;;) {
        unsigned t = 0;
        if (x.lane == 0) t = atomicAdd(ticket, 1u);
        t = __shfl_sync(0xFFFFFFFFu, t, 0);
        if (t >= nwork) return;
        const uint32_t c = work[t];
        if (reset) {
            if (x.lane == 0)
                for (uint32_t at = chain_start[c]; at < chain_start[c + 1]; at++) {
                    const Read &r = reads[chain_reads[at]];
                    heads[r.fkey].created_by = -1; heads[r.rkey].created_by = -1;
                }
            __syncwarp();
        }
        for (uint32_t at = chain_start[c]; at < chain_start[c + 1]; at++) {
            const uint32_t i = chain_reads[at];
            const int e = process_read(x, reads[i], (int)i, heads, bseqs, bquals, fwd, qfwd, rev, qrev, s);
            if (e) {
                if (redo_list) {                       // too big for this geometry (or an error the second launch will name)
                    if (x.lane == 0) redo_list[atomicAdd(redo_count, 1u)] = c;
                    break;
                }
                if (x.lane == 0 && atomicCAS(&err[0], 0, e) == 0) err[1] = (int)i;   // the whole call fails: remember the first reason
                return;
            }
        }
    }
}

__global__ void k_merge_init(BlockHead *heads, uint32_t nkeys)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nkeys) { BlockHead h; memset(&h, 0, sizeof h); h.created_by = -1; heads[k] = h; }
}

// the reverse strand of every read (ReverseComplementSequence, fastq_seq.c:153-176): a warp per read
__global__ void k_merge_strands(const Read *reads, uint32_t nreads, const char *fwd, const char *qfwd, char *rev, char *qrev)
{
    const uint32_t r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= nreads) return;
    const uint32_t off = reads[r].off, L = reads[r].slen;
    for (uint32_t i = threadIdx.x & 31; i < L; i += 32) {
        const unsigned char c = (unsigned char)fwd[off + L - 1 - i], u = c & 0xDF, low = c & 0x20;
        // the IUPAC complement in either case, a blank for everything else (the table of fastq_seq.c:146-150)
        unsigned char o = ' ';
        if (c >= 64 && c < 128) {
            switch (u) {
            case 'A': o = 'T'; break; case 'T': o = 'A'; break; case 'C': o = 'G'; break; case 'G': o = 'C'; break;
            case 'B': o = 'V'; break; case 'V': o = 'B'; break; case 'D': o = 'H'; break; case 'H': o = 'D'; break;
            case 'K': o = 'M'; break; case 'M': o = 'K'; break; case 'R': o = 'Y'; break; case 'Y': o = 'R'; break;
            case 'N': o = 'N'; break; case 'S': o = 'S'; break; case 'W': o = 'W'; break; case 'X': o = 'X'; break;
            default: break;
            }
            if (o != ' ') o |= low;
        }
        rev[off + i] = (char)o;
        qrev[off + i] = qfwd[off + L - 1 - i];
    }
    if ((threadIdx.x & 31) == 0) { rev[off + L] = 0; qrev[off + L] = 0; }
}

// what the host needs of EVERY key: who created its block (the order of the reference's map) and whether it is printed
__global__ void k_merge_verdicts(const BlockHead *heads, uint32_t nkeys, uint32_t lo, uint32_t hi, int include_all, int32_t *created_by, uint8_t *print)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nkeys) return;
    const BlockHead h = heads[k];
    created_by[k] = h.created_by;
    print[k] = h.created_by >= 0 && block_verdict(h, lo, hi, include_all != 0).print;
}

// the blocks that will be printed, side by side
__global__ void k_merge_gather(const uint32_t *which, uint32_t n, const BlockHead *heads, const char *bseqs, const char *bquals, BlockHead *ohead,
                               char *oseq, char *oqual)
{
    const uint32_t b = blockIdx.x;
    if (b >= n) return;
    const size_t from = (size_t)which[b] * MAXB, to = (size_t)b * MAXB;
    for (int a = threadIdx.x; a < MAXB; a += blockDim.x) { oseq[to + a] = bseqs[from + a]; oqual[to + a] = bquals[from + a]; }
    if (threadIdx.x == 0) ohead[b] = heads[which[b]];
}

static const char *err_text(int e)
{
    switch (e) {
    case ERR_FLANK_TOO_LONG: return "a flank longer than 383 bases (the alignment matrix is held in shared memory)";
    case ERR_BLOCK_TOO_LONG: return "a block longer than 799 bases";
    case ERR_TOO_MANY_ALLELES: return "more than 24 different copy numbers in one block";
    }
    return "?";
}

}  // namespace bmerge

using namespace bmerge;

namespace {
struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    template <class T> T *as() { return (T *)p; }
};
}

extern "C" int bstr_merge_reads(bstr_ctx *ctx, const uint8_t *fastq, uint64_t nbytes, int klength, int min_threshold, int max_threshold,
                                int include_all, int out_fd, uint64_t stats[6])
{
    using bstr::set_error;
    using bstr::cuda_fail;
    if (!ctx || (!fastq && nbytes)) return BSTR_ERR_PARAM;
    if (klength < 1 || klength > 1000) { set_error(ctx, "bstr_merge_reads: klength %d", klength); return BSTR_ERR_PARAM; }
    if (klength % 2 == 0) klength--;                                       // merge_STR_reads.c:557-560
    if (include_all && min_threshold < 2) {
        set_error(ctx, "bstr_merge_reads: --all with min_threshold < 2: the reference stops on its own assertion at the first unmerged block (merge_STR_reads.c:468)");
        return BSTR_ERR_PARAM;
    }
    const auto t0 = std::chrono::steady_clock::now();
    Plan P;
    if (!make_plan(fastq, nbytes, klength, P, 0, false)) { set_error(ctx, "bstr_merge_reads: %s", P.error.c_str()); return BSTR_ERR_PARAM; }
    const uint32_t nreads = (uint32_t)P.reads.size(), nkeys = (uint32_t)P.keys.size(), nchains = (uint32_t)P.chain_start.size() - 1;
    std::string out;
    uint64_t st_blocks = 0, st_printed = 0, st_longest = 0, st_redone = 0;
    double ms_plan = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(), ms_dev = 0, ms_kernel = 0;
    if (nreads) {
        // longest chains first
        std::vector<uint32_t> order(nchains);
        std::iota(order.begin(), order.end(), 0u);
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
            return P.chain_start[a + 1] - P.chain_start[a] > P.chain_start[b + 1] - P.chain_start[b];
        });
        st_longest = P.chain_start[order[0] + 1] - P.chain_start[order[0]];
        const auto t1 = std::chrono::steady_clock::now();
        CUDA_TRY(cudaSetDevice(ctx->dev));
        cudaStream_t st = ctx->stream;
        DevBuf d_reads, d_cs, d_cr, d_co, d_heads, d_bseq, d_bqual, d_fwd, d_qfwd, d_rev, d_qrev, d_misc, d_which, d_oseq, d_oqual, d_made, d_print, d_ohead;
        const size_t blob = P.fwd.size();
#define BM_UP(buf, src, bytes)                                                                                     \
        do {                                                                                                           \
            if (cudaMalloc(&buf.p, (bytes) ? (bytes) : 1) != cudaSuccess) return cuda_fail(ctx, cudaGetLastError(), "merge: cudaMalloc", __LINE__); \
            if ((src) && (bytes)) CUDA_TRY(cudaMemcpyAsync(buf.p, (src), (bytes), cudaMemcpyHostToDevice, st));          \
        } while (0)
        BM_UP(d_reads, P.reads.data(), (size_t)nreads * sizeof(Read));
        BM_UP(d_cs, P.chain_start.data(), (size_t)(nchains + 1) * 4);
        BM_UP(d_cr, P.chain_reads.data(), (size_t)nreads * 4);
        BM_UP(d_co, order.data(), (size_t)nchains * 4);
        BM_UP(d_fwd, P.fwd.data(), blob);
        BM_UP(d_qfwd, P.qfwd.data(), blob);
        BM_UP(d_rev, (const void *)nullptr, blob);
        BM_UP(d_qrev, (const void *)nullptr, blob);
        BM_UP(d_made, (const void *)nullptr, (size_t)nkeys * 4);
        BM_UP(d_print, (const void *)nullptr, (size_t)nkeys);
        BM_UP(d_heads, (const void *)nullptr, (size_t)nkeys * sizeof(BlockHead));
        BM_UP(d_bseq, (const void *)nullptr, (size_t)nkeys * MAXB);
        BM_UP(d_bqual, (const void *)nullptr, (size_t)nkeys * MAXB);
        BM_UP(d_misc, (const void *)nullptr, 32);                       // ticket, error code, error read, chains to redo, ticket of the second launch
        CUDA_TRY(cudaMemsetAsync(d_misc.p, 0, 32, st));
        k_merge_init<<<(nkeys + 255) / 256, 256, 0, st>>>(d_heads.as<BlockHead>(), nkeys);
        k_merge_strands<<<(nreads + 7) / 8, 256, 0, st>>>(d_reads.as<Read>(), nreads, d_fwd.as<char>(), d_qfwd.as<char>(), d_rev.as<char>(), d_qrev.as<char>());
        // geometry: sized for the longest flank of the input with room to grow; chains that outgrow it are redone
        int longest_flank = 0;
        for (const Read &r : P.reads)
            longest_flank = std::max(longest_flank, std::max(std::max((int)r.fzstart, (int)r.slen - (int)r.fend), std::max((int)r.rzstart, (int)r.slen - (int)r.rend)));
        const char *force = getenv("BSTR_MERGE_MAXF");                  // tests: a first geometry that most chains outgrow
        const Geometry small = geometry_for(force ? atoi(force) : longest_flank + longest_flank / 4 + 16), full = geometry_for(MAXF);
        const bool two_pass = small.maxf < full.maxf;
        DevBuf d_redo;
        if (two_pass) BM_UP(d_redo, (const void *)nullptr, (size_t)nchains * 4);
        CUDA_TRY(cudaFuncSetAttribute(k_merge_chains, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(full.warps * full.warp_bytes > small.warps * small.warp_bytes
                                                                                                            ? full.warps * full.warp_bytes : small.warps * small.warp_bytes)));
        cudaEvent_t e0, e1;
        CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1));
        auto launch = [&](const Geometry &g, const uint32_t *work, uint32_t nwork, unsigned *ticket, uint32_t *redo, unsigned *redo_count, int reset) {
            const unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count * g.ctas_per_sm, ((uint64_t)nwork + g.warps - 1) / g.warps);
            k_merge_chains<<<grid ? grid : 1, g.warps * 32, (size_t)g.warps * g.warp_bytes, st>>>(
                d_reads.as<Read>(), d_cs.as<uint32_t>(), d_cr.as<uint32_t>(), work, nwork, d_heads.as<BlockHead>(), d_bseq.as<char>(), d_bqual.as<char>(),
                d_fwd.as<char>(), d_qfwd.as<char>(), d_rev.as<char>(), d_qrev.as<char>(), ticket, d_misc.as<int>() + 1, redo, redo_count, reset, g);
        };
        cudaEventRecord(e0, st);
        unsigned redone = 0;
        if (!two_pass) {
            launch(full, d_co.as<uint32_t>(), nchains, d_misc.as<unsigned>(), nullptr, nullptr, 0);
        } else {
            launch(small, d_co.as<uint32_t>(), nchains, d_misc.as<unsigned>(), d_redo.as<uint32_t>(), d_misc.as<unsigned>() + 3, 0);
            if (cudaMemcpyAsync(&redone, d_misc.as<unsigned>() + 3, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
                cudaEventDestroy(e0); cudaEventDestroy(e1);
                return cuda_fail(ctx, cudaGetLastError(), "k_merge_chains", __LINE__);
            }
            if (redone) launch(full, d_redo.as<uint32_t>(), redone, d_misc.as<unsigned>() + 4, nullptr, nullptr, 1);
        }
        cudaEventRecord(e1, st);
        st_redone = redone;
        cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return cuda_fail(ctx, le, "k_merge_chains", __LINE__); }
        k_merge_verdicts<<<(nkeys + 255) / 256, 256, 0, st>>>(d_heads.as<BlockHead>(), nkeys, (uint32_t)min_threshold, (uint32_t)max_threshold, include_all,
                                                              d_made.as<int32_t>(), d_print.as<uint8_t>());
        std::unique_ptr<int32_t[]> made(new int32_t[nkeys]);
        std::unique_ptr<uint8_t[]> printed(new uint8_t[nkeys]);
        int misc[4] = {0, 0, 0, 0};
        cudaError_t ce = cudaMemcpyAsync(made.get(), d_made.p, (size_t)nkeys * 4, cudaMemcpyDeviceToHost, st);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(printed.get(), d_print.p, (size_t)nkeys, cudaMemcpyDeviceToHost, st);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(misc, d_misc.p, 16, cudaMemcpyDeviceToHost, st);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
        float kms = 0;
        if (ce == cudaSuccess) cudaEventElapsedTime(&kms, e0, e1);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (ce != cudaSuccess) return cuda_fail(ctx, ce, "merge: kernel / download", __LINE__);
        ms_kernel = kms;
        if (misc[1]) {
            set_error(ctx, "bstr_merge_reads: read %d: %s", misc[2] + 1, err_text(misc[1]));
            return BSTR_ERR_TOO_LARGE;
        }
        // the keys in the order their blocks entered the reference's map = the order of the reads that made them
        // (a read makes at most one block)
        std::vector<uint32_t> by_read(nreads, 0xFFFFFFFFu);
        for (uint32_t k = 0; k < nkeys; k++) if (made[k] >= 0) by_read[(uint32_t)made[k]] = k;
        std::vector<uint32_t> inserted;
        inserted.reserve(nkeys);
        for (uint32_t r = 0; r < nreads; r++) if (by_read[r] != 0xFFFFFFFFu) inserted.push_back(by_read[r]);
        st_blocks = inserted.size();
        const std::vector<uint32_t> listed = table_order(P.keys, inserted);
        std::vector<uint32_t> which;
        for (uint32_t k : listed) if (printed[k]) which.push_back(k);
        st_printed = which.size();
        std::vector<char> oseq(which.size() * (size_t)MAXB), oqual(which.size() * (size_t)MAXB);
        std::vector<BlockHead> heads(which.size());
        if (!which.empty()) {
            BM_UP(d_which, which.data(), which.size() * 4);
            BM_UP(d_oseq, (const void *)nullptr, oseq.size());
            BM_UP(d_oqual, (const void *)nullptr, oqual.size());
            BM_UP(d_ohead, (const void *)nullptr, which.size() * sizeof(BlockHead));
            k_merge_gather<<<(unsigned)which.size(), 128, 0, st>>>(d_which.as<uint32_t>(), (uint32_t)which.size(), d_heads.as<BlockHead>(), d_bseq.as<char>(),
                                                                  d_bqual.as<char>(), d_ohead.as<BlockHead>(), d_oseq.as<char>(), d_oqual.as<char>());
            CUDA_TRY(cudaGetLastError());
            CUDA_TRY(cudaMemcpyAsync(oseq.data(), d_oseq.p, oseq.size(), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(oqual.data(), d_oqual.p, oqual.size(), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(heads.data(), d_ohead.p, which.size() * sizeof(BlockHead), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
        }
#undef BM_UP
        unsigned bindex = 1;
        for (size_t b = 0; b < which.size(); b++)
            print_block(out, bindex, P.keys.key(which[b]), heads[b], &oseq[b * (size_t)MAXB], &oqual[b * (size_t)MAXB], (unsigned)min_threshold,
                        (unsigned)max_threshold, include_all != 0);
        ms_dev = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
    }
    if (out_fd >= 0) {
        size_t done = 0;
        while (done < out.size()) {
            const ssize_t n = write(out_fd, out.data() + done, out.size() - done);
            if (n <= 0) { set_error(ctx, "bstr_merge_reads: write failed"); return BSTR_ERR_IO; }
            done += (size_t)n;
        }
    }
    if (stats) { stats[0] = nreads; stats[1] = nkeys; stats[2] = nchains; stats[3] = st_blocks; stats[4] = st_printed; stats[5] = st_longest; }
    if (getenv("BSTR_TRACE"))
        fprintf(stderr, "[bstr_merge] %u reads, %u keys, %u chains (longest %llu), %llu blocks, %llu printed: plan %.1f ms, device part %.1f ms (kernels %.2f ms, %llu chains redone with the largest geometry)\n",
                nreads, nkeys, nchains, (unsigned long long)st_longest, (unsigned long long)st_blocks, (unsigned long long)st_printed, ms_plan, ms_dev, ms_kernel,
                (unsigned long long)st_redone);
    return BSTR_OK;
}
