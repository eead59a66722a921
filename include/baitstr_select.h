/*
 * baitstr_select.h -- C ABI of libbaitstr_select.so, the B200 (sm_100a) implementation of
 * BaitSTR's per-read STR detection stage (select_STR_reads).
 *
 * This is the drop-in boundary: everything the reference's Python entry point
 *     select_str_reads(filenames, num_minimum_copies, flanking_distance, illumina_quals,
 *                      check_strs, only_k_mers, remove_homopolymers, debug_flag, altalgo,
 *                      num_processes, chunksize)        /root/reference/src/select_STR_reads.pyx:339-373
 * does between "open the FASTQ files" and "print the annotated FASTQ" happens behind
 * these calls.  Plain C types only; no CUDA, PyTorch or C++ types cross the boundary.
 * There is no CPU fallback: without a usable CUDA device bstr_create() fails.
 *
 * Threading: a context is used by one caller thread at a time.  Buffers returned through
 * bstr_batch_out belong to the context and stay valid until the next call on it.
 */
#ifndef BAITSTR_SELECT_H
#define BAITSTR_SELECT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BSTR_ABI_VERSION 2

/* error codes (negative return values) */
enum {
    BSTR_OK = 0,
    BSTR_ERR_NO_DEVICE = -1,    /* no CUDA device / driver: the product path cannot run      */
    BSTR_ERR_CUDA = -2,         /* a CUDA call failed; bstr_last_error() has the text        */
    BSTR_ERR_PARAM = -3,        /* bad argument (e.g. numcopies < 0, no period selected)     */
    BSTR_ERR_IO = -4,           /* open/read/inflate/write failed                            */
    BSTR_ERR_FASTQ_HEADER = -10,    /* record does not start with '@'   (fastq.pyx:77)       */
    BSTR_ERR_FASTQ_BASES = -11,     /* base outside acgtnACGTN.         (fastq.pyx:62-67,81) */
    BSTR_ERR_FASTQ_SEPARATOR = -12, /* third line does not start '+'    (fastq.pyx:86)       */
    BSTR_ERR_FASTQ_TRUNCATED = -13, /* file ends inside a record        (fastq.pyx:86)       */
    BSTR_ERR_TOO_LARGE = -14    /* a single record does not fit the batch buffer             */
};

/*
 * Scan parameters = the reference's command-line options
 * (/root/reference/src/select_STR_reads:74-90,97-123).
 */
typedef struct bstr_params {
    int32_t abi_version;     /* BSTR_ABI_VERSION                                            */
    int32_t num_copies;      /* -n/--numcopies           default 2                           */
    int32_t flank;           /* -f/--flanks              default 31                          */
    int32_t kmask;           /* bit k (2..6) set => -k/--kmer given; 0 => all of 2..6        */
    int32_t remove_homopolymers; /* -r/--removehm                                           */
    int32_t altalgo;         /* 1 = default algorithm; 0 = -x/--altalgo given                */
    int32_t illumina_quals;  /* -i/--illumina: Phred+64 in, Phred+33 out                     */
    int32_t n_devices;       /* GPUs to use; 0 => all visible                                */
    int32_t devices[16];     /* CUDA ordinals when n_devices > 0                             */
    int32_t host_threads;    /* reader/inflate threads hint (from -p); 0 => default          */
    int32_t want_results;    /* bstr_scan_batch also returns the 16-byte per-read records    */
    int32_t reserved0;
    uint64_t batch_bytes;    /* raw FASTQ bytes per device batch; 0 => default (8 MiB)       */
} bstr_params;

/* 16 bytes per input record, in input order. fck == 0 <=> read not selected.
 * Coordinates as printed by PrintSequence (select_STR_reads.pyx:174-178): fstart in the
 * read as given, rstart in the reverse-complemented read. */
typedef struct bstr_read_result {
    uint32_t fstart;
    uint32_t rstart;
    uint32_t fck;            /* forward copies << 3 | period */
    uint32_t rck;            /* reverse-complement copies << 3 | period */
} bstr_read_result;

typedef struct bstr_batch_out {
    const uint8_t *out;      /* annotated FASTQ of the selected reads, input order           */
    uint64_t out_bytes;
    uint64_t consumed;       /* input bytes covered by whole records (== nbytes if final)    */
    uint64_t n_records;
    uint64_t n_selected;
    const bstr_read_result *results;   /* NULL unless params.want_results                    */
    int64_t bad_record;      /* on a BSTR_ERR_FASTQ_* error: 0-based record index in batch   */
} bstr_batch_out;

typedef struct bstr_stats {
    uint64_t records, selected, in_bytes, out_bytes, batches;
    uint64_t fallback_batches;                     /* batches re-run through the generic kernels */
    double h2d_ms, kernel_ms, d2h_ms, wall_ms;     /* summed over batches (CUDA events)      */
    uint64_t device_batches[16];                   /* batches collected per entry of params.devices */
} bstr_stats;

/* result of the device-resident benchmark loop */
typedef struct bstr_bench_out {
    double ms_total;         /* CUDA-event time of `steps` passes on the launch stream       */
    double ms_kernel[8];     /* per-stage device time per step (all batches), ms             */
    char kernel_name[8][32];
    int32_t n_kernels;
    int32_t launches_per_step;
    uint64_t n_records, n_selected, in_bytes, out_bytes;
    uint64_t out_checksum;   /* FNV-1a of the output bytes of the last pass                  */
} bstr_bench_out;

typedef struct bstr_ctx bstr_ctx;

/* Create a context on the device(s) named in p.  On failure returns NULL and writes a
 * message to err (if errlen > 0).  Replaces Pool(num_processes) + pattern compilation
 * (select_STR_reads.pyx:347-357). */
bstr_ctx *bstr_create(const bstr_params *p, char *err, size_t errlen);
void bstr_destroy(bstr_ctx *ctx);
const char *bstr_last_error(const bstr_ctx *ctx);

/* Whole stage for a list of FASTQ files (".gz" by file-name suffix, fastq.pyx:50), in argv
 * order, selected reads written to out_fd, "Done processing <path>\n" to err_fd after each
 * file.  Replaces the file loop of select_STR_reads.pyx:359-370.  Returns BSTR_OK or <0. */
int bstr_run_files(bstr_ctx *ctx, const char *const *paths, int npaths, int out_fd, int err_fd);

/* One batch of raw FASTQ bytes in host memory, starting at a record boundary.  Processes
 * every record that is complete in the buffer (all of it when `final`, with end-of-file
 * semantics of fastq.pyx:72-94 for the last line) and returns the formatted output.
 * H2D copy, kernels and D2H copy are all inside this call. */
int bstr_scan_batch(bstr_ctx *ctx, const uint8_t *raw, uint64_t nbytes, int final,
                    bstr_batch_out *out);

/* The same work as bstr_scan_batch, split in two so that several batches are in flight:
 * "pinned double-buffered H2D/D2H batches on CUDA streams".  A context has bstr_lanes()
 * lanes (3 per device; lane l runs on device l % n_devices).  bstr_submit_batch enqueues the
 * H2D copy, the kernels and the state read-back on the lane's stream and returns at once;
 * bstr_collect_batch waits for that batch and copies back exactly the bytes produced.  `raw`
 * must stay valid (and should be page-locked, bstr_host_alloc) until the collect; the output
 * stays valid until the next submit on the same lane.  Replaces the blocking
 * pool.map(func, chunk) of select_STR_reads.pyx:363; the caller keeps batch order. */
int bstr_lanes(bstr_ctx *ctx);
int bstr_submit_batch(bstr_ctx *ctx, int lane, const uint8_t *raw, uint64_t nbytes, int final);
int bstr_collect_batch(bstr_ctx *ctx, int lane, bstr_batch_out *out);

/* Per-read matcher hook (the GPU counterpart of ProcessSequence / MatchPatternsAndReportBest,
 * select_STR_reads.pyx:106,294): nreads sequences concatenated in `seqs`, read i occupying
 * [offsets[i], offsets[i+1]).  Bases may be lower case; they are not validated here. */
int bstr_match_reads(bstr_ctx *ctx, const uint8_t *seqs, const uint64_t *offsets,
                     uint64_t nreads, bstr_read_result *results);

/* -d/--debug: what the reference prints per read when debug_flag is set
 * (select_STR_reads.pyx:296-311 with :76-87 / :138-140 and :157-163).  For read i the codes of
 * the regex hits the reference rejects with a message -- 'N' "N in motif.", '5' "Insufficient
 * 5' flank.", '3' "Insufficient 3' flank.", in the reference's order -- are written to
 * codes[code_offsets[i] ...) for the forward strand and to the second half of the slot for the
 * reverse complement; every slot must hold 2 * bstr_debug_codes_cap(ctx, read length) bytes.
 * info[8 i ..] = {forward: number of codes, 1 if the strand has no best match ("0 valid
 * matches."), tail code and value of the -x algorithm's message about its winner ('5' / '3',
 * "Insufficient 5' flank.: <value>"), then the same four for the reverse strand, which the
 * reference only looks at when the forward strand has a match}.  Sequences as for
 * bstr_match_reads. */
int bstr_debug_reads(bstr_ctx *ctx, const uint8_t *seqs, const uint64_t *offsets, uint64_t nreads,
                     const uint64_t *code_offsets, uint8_t *codes, uint32_t *info);
uint32_t bstr_debug_codes_cap(const bstr_ctx *ctx, uint32_t read_length);

/* Upload `nbatches` batches (each a whole number of records, as bstr_scan_batch with
 * final = 1) and run the device pipeline over all of them `warmup + steps` times with the
 * input resident in HBM (no host copies inside the timed region); optionally evict L2
 * between passes.  Used by bench.py for the kernel-only throughput and the roofline. */
int bstr_bench_resident(bstr_ctx *ctx, const uint8_t *const *raws, const uint64_t *nbytes,
                        int nbatches, int steps, int warmup, int flush_l2, bstr_bench_out *out);

/* Copy-only probe of the host<->device path that the lanes use (no kernels): the jobs
 * (raws[i], nbytes[i]) are dealt to the context's lanes exactly as bench.py's end-to-end leg
 * deals its sub-batches (lane i % bstr_lanes(), at most one job per lane in flight, wait for
 * the oldest before re-using its lane).  mode bit 0: copy the input host -> device; bit 1:
 * copy out_bytes[i] bytes device -> page-locked host (after the lane's H2D completed, as
 * bstr_collect_batch does).  One cudaMemcpyAsync per copy.  *ms = host wall time of `reps`
 * passes.  Names the ceiling of the end-to-end numbers; no counterpart in the reference. */
int bstr_pcie_probe(bstr_ctx *ctx, const uint8_t *const *raws, const uint64_t *nbytes,
                    const uint64_t *out_bytes, int njobs, int mode, int reps, double *ms);

/* bstr_scan_batch for callers that may not allocate device memory (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  This entry point has no counterpart in
 * the reference.  It does the work of
 * bstr_scan_batch -- H2D copy, the fused tile kernel, D2H copy -- and calls no cudaMalloc, cudaFree,
 * cudaHostAlloc or other allocating routine: every device byte it needs (input staging, output,
 * look-back words, batch state) lives in ONE static device array of about 12.1 MiB inside the
 * library, which every GPU that loads the library holds until the process exits.
 *   raw, nbytes    host memory; nbytes <= BSTR_NOALLOC_MAX_BYTES, else BSTR_ERR_TOO_LARGE
 *   out            caller's buffer (host or device) of out_capacity bytes; receives the annotated
 *                  FASTQ.  If it is too small the call fails with BSTR_ERR_TOO_LARGE and
 *                  info->out_bytes says how much is needed (never more than 2 x nbytes + 4096 here;
 *                  beyond that: BSTR_ERR_TOO_LARGE as well)
 *   info           counts as for bstr_scan_batch; info->out == out, info->results == NULL
 * Inputs the fused kernel's tile geometry cannot hold (a record that runs more than 4 KiB past the
 * end of the ~40 KiB tile it starts in, ...) are
 * refused with BSTR_ERR_TOO_LARGE instead of being re-run through the generic kernels, which need
 * allocated tables.  All calls on a device share the array and it keeps the previous call's
 * contents; what must start at zero is cleared at the start of each call.  Calls that overlapped
 * on the GPU would race on it: within a process the entry point serialises them with a lock held
 * until the call's last synchronize (other processes have their own copy of the array). */
#define BSTR_NOALLOC_MAX_BYTES (4u << 20)
int bstr_scan_batch_noalloc(bstr_ctx *ctx, const uint8_t *raw, uint64_t nbytes, int final,
                            uint8_t *out, uint64_t out_capacity, bstr_batch_out *info);

/* Page-locked host memory for callers that want their batches DMA'd without a staging
 * copy (bench.py's end-to-end leg).  Plain malloc'd buffers work too, only slower. */
void *bstr_host_alloc(size_t nbytes);
void bstr_host_free(void *p);

void bstr_get_stats(const bstr_ctx *ctx, bstr_stats *st);
void bstr_reset_stats(bstr_ctx *ctx);

/* Host-only test hook (no GPU needed): the tile geometry of the fused kernel (csrc/fused.cuh:
 * TileGeom) for batches whose records average record_bytes bytes, for the instantiation with
 * window_words = 2, 4, 6 or 8.  out = {tile bytes, CTA threads, window bytes, 16-byte chunks,
 * chunks per thread, record slots, staging bytes, dynamic shared memory}.  No counterpart in
 * the reference (its chunk size is the -s option, src/select_STR_reads.pyx:321-337). */
int bstr_geometry_probe(double record_bytes, int window_words, uint32_t out[8]);

/* Host-only test hook for the ingest side (no GPU needed): inflate the gzip file `path`
 * (single- or multi-member, bgzip included) on `threads` threads with the parallel inflater
 * that bstr_run_files uses for "*.gz" inputs (the reference inflates on one thread through
 * Python's gzip module, src/fastq.pyx:50-51) and write the bytes to out_fd (-1: discard).
 * range_bytes: compressed bytes per decoder work item; seg_bytes: inflated bytes per output
 * segment.  stats6 (may be NULL): in bytes, out bytes, ranges searched, runs accepted, runs
 * decoded again from the true position, gzip members.  0, or BSTR_ERR_IO with a message in err
 * (not gzip, corrupt deflate data, CRC-32/ISIZE mismatch, truncated file, bytes after the last
 * member that are neither zero padding nor another member -- Python's gzip module rejects those
 * too).  threads <= 0 selects the serial zlib-based reader that bstr_run_files uses for small
 * files; both readers accept and reject the same files. */
int bstr_gunzip(const char *path, int out_fd, int threads, uint64_t range_bytes, uint64_t seg_bytes,
                uint64_t *stats6, char *err, size_t errlen);

/* The same for the GPU inflater (csrc/ginflate.cu), which bstr_run_files uses for "*.gz" inputs of
 * some size: the compressed file is cut into windows of window_bytes (0: 256 MiB) and ranges of
 * range_bytes (0: 128 KiB); one warp per range decodes DEFLATE into 16-bit symbols from a guessed
 * block start, a chain check accepts the runs, the 32 KiB history is propagated, symbols become
 * bytes and every member's CRC-32 / ISIZE is checked -- all on CUDA device `device`; only the
 * inflated bytes this hook writes to out_fd (-1: discard) come back to the host.  Accepts and rejects
 * the same files as bstr_gunzip / Python's gzip module (the reference reads through it,
 * src/fastq.pyx:50-51).  stats6 as for bstr_gunzip.  Needs a CUDA device. */
int bstr_gunzip_gpu(const char *path, int out_fd, int device, uint64_t window_bytes, uint64_t range_bytes,
                    uint64_t *stats6, char *err, size_t errlen);

/* First slice of the next stage: merge_STR_reads on the GPU.  Replaces
 * MergeShortTandemRepeatReads, /root/reference/src/merge_STR_reads.c:308-509 (with Align :43-213 and
 * AlignFlanks :215-306): `fastq` is the select stage's output (annotated names), the polymorphic
 * blocks are written to out_fd byte for byte as `merge_STR_reads [--min_threshold=..]
 * [--max_threshold=..] [--all] klength reads.str.fq` prints them -- same blocks, same consensus
 * sequences and qualities, same order and numbering (the order is that of the reference's hash map,
 * replayed).  The reads are cut into chains that cannot influence each other and a warp walks each
 * chain; the local alignments run on the 32 lanes.
 *   klength        as the reference's: an even value is lowered by one (:557-560)
 *   min_threshold / max_threshold / include_all   the options of the same names (defaults 4, 10000, 0)
 *   out_fd         file descriptor for the blocks, < 0 = none
 *   stats          NULL or {reads, keys, chains, blocks, blocks printed, longest chain}
 * Refused with an error instead of reproducing undefined behaviour of the reference: a flank shorter
 * than klength (it reads outside the read), --all with min_threshold < 2 (its assertion :468 stops
 * it), flanks > 383 or blocks > 799 bases. */
int bstr_merge_reads(bstr_ctx *ctx, const uint8_t *fastq, uint64_t nbytes, int klength, int min_threshold,
                     int max_threshold, int include_all, int out_fd, uint64_t stats[6]);

/* First slice of the stage after next: the k-mer counting pass of
 * extend_STR_reads on the GPU.  Replaces ReadAndCountNonSingletonKmers,
 * /root/reference/src/extend_STR_reads.c:31-194 (Bloom filter of singletons + sparse hash map of
 * canonical 2-bit k-mers, kmer.c:4-47, then the threshold pruning :187-193): every k-mer of every
 * read is counted exactly in one pass in an open-addressing table in HBM, and bstr_kmers_finish returns
 * { canonical k-mer : min(count, 255) } for min_threshold <= count <= max_threshold, sorted by k-mer.
 * The k-mer arithmetic is the reference's bit for bit (uint64_t, klen in 2..31), including what a base
 * outside ACGTacgt does to the rolling word (N and n read as A, anything else adds -1:
 * utilities.h:61-75).  min_threshold must be >= 2: below that the reference's own result depends on
 * its Bloom filter's false positives (seeded from time()).
 *   create     expected_kmers = distinct k-mers the table must hold, singletons included
 *   add_fastq  raw FASTQ in host memory: whole 4-line records, newline-terminated; any number of calls
 *   add_resident  see below
 *   finish     the arrays belong to the handle and stay valid until it is destroyed
 *   stats      out4 = {records, k-mers counted, distinct k-mers, table slots} */
typedef struct bstr_kmers bstr_kmers;
bstr_kmers *bstr_kmers_create(bstr_ctx *ctx, int klen, uint64_t expected_kmers);
int bstr_kmers_add_fastq(bstr_kmers *k, const uint8_t *raw, uint64_t nbytes);
/* The same for the batch the select stage has just scanned: bstr_scan_batch leaves its input in device
 * memory, and the first nbytes of it (<= the `consumed` it reported: whole records) are counted where
 * they lie -- the reads cross PCIe once for both stages.  The handle must belong to the context that
 * ran bstr_scan_batch; no other batch call in between. */
int bstr_kmers_add_resident(bstr_kmers *k, uint64_t nbytes);
int bstr_kmers_finish(bstr_kmers *k, int min_threshold, int max_threshold, const uint64_t **kmers,
                      const uint8_t **counts, uint64_t *n);
void bstr_kmers_stats(const bstr_kmers *k, uint64_t out4[4]);
/* The rest of extend_STR_reads over the same table: the walk through the k-mer graph on both sides of
 * every merged block and the contigs.  Replaces the block loop of ExtendShortTandemRepeatReads,
 * /root/reference/src/extend_STR_reads.c:909-998, with ExtendBackward / ExtendForward :471-734,
 * CheckForSNPBackwards / Forwards :218-438, FindFirstGoodKmer :809-850, PercentIdentity :740-772 and
 * PrintContig :776-807.  `blocks_fastq` is the merge stage's output; the contigs are written to
 * out_fd byte for byte as the reference prints them.  A warp walks each block, eight lanes probing
 * the eight neighbours of a k-mer at once.  As in the reference, the table must have been counted
 * over the merged blocks' file as well as the read files (it passes argv[4 ..] to the counting pass,
 * :890-897): call bstr_kmers_add_fastq with the blocks first.
 *   min_threshold / max_threshold   k-mers outside are not in the table (defaults 2, 255); min >= 2
 *   flanks                          the --flanks option (default 1024), 2..2048 here
 *   stats                           NULL or {blocks, contigs printed, ends not extended, k-mers within the thresholds}
 * A block whose right flank is shorter than klen is skipped (the reference walks on from the bytes
 * behind the block's text there). */
int bstr_kmers_extend(bstr_kmers *k, int min_threshold, int max_threshold, int flanks, const uint8_t *blocks_fastq,
                      uint64_t nbytes, int out_fd, uint64_t stats[4]);
void bstr_kmers_destroy(bstr_kmers *k);

/* number of CUDA devices visible (0 if none / no driver) */
int bstr_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* BAITSTR_SELECT_H */
