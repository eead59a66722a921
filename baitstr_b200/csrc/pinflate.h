// Parallel inflate of ONE gzip file for the ingest side of bstr_run_files.
//
// The reference opens "*.gz" with Python's gzip module and inflates on one thread
// (src/fastq.pyx:50-51).  Here the compressed file is cut into fixed ranges; a pool of threads
// decodes the ranges at the same time, each from the first deflate block boundary it can find in
// its range, WITHOUT knowing the 32 KiB of history before it: back-references that reach into
// the unknown history are written as 16-bit markers and are replaced by real bytes once the
// range before it is complete (the approach of pugz / rapidgzip, restated from their papers).
// A range is accepted only if the range before it ended on exactly the bit it started on;
// otherwise it is decoded again from the true position, so the result never depends on the
// block search being right.  CRC-32 and ISIZE of every gzip member are verified.
//
// Consumers see the inflated stream as fixed-size SEGMENTS that can be claimed and filled from
// several threads (the same shape as the parallel reader for plain files in stream_host.cu).
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <vector>

namespace pinflate {

struct SegmentInfo {                  // what fill() learned about one segment
    uint64_t newlines = 0;
    struct Part { uint64_t len; uint32_t crc; };
    std::vector<Part> parts;          // the segment split at gzip member ends, CRC-32 of each part
};

struct Stats {
    uint64_t in_bytes = 0, out_bytes = 0, ranges = 0, accepted = 0, redone = 0, members = 0;
    double decode_ms = 0, search_ms = 0, redo_ms = 0, wait_ms = 0;
};

class Engine;                          // pinflate.cpp

class Stream {
public:
    Stream();
    ~Stream();
    // data[0, size) is the whole .gz file (kept mapped by the caller until close()).
    // threads: decoder threads; range_bytes: compressed bytes per range; seg_bytes: inflated bytes per segment;
    // pending_segments: how far (in segments) decoding may run ahead of fill() - bounds the memory held.
    // false: not a gzip file (error() says why).
    bool open(const uint8_t *data, size_t size, int threads, size_t range_bytes, size_t seg_bytes, int pending_segments);
    // Blocks until the next segment of the inflated stream is known.  Returns its index (0, 1, ...),
    // -1 after the last one, -2 on error.  Thread-safe; indices are handed out in order.
    long claim(uint64_t *off, size_t *len, bool *last);
    // Write the bytes of a claimed segment to dst[0, len).  Thread-safe, any order.
    bool fill(uint64_t off, size_t len, uint8_t *dst, SegmentInfo *info);
    // Verify member CRC/ISIZE; call once per segment IN ORDER with fill()'s info.  false: mismatch.
    bool fold(const SegmentInfo &info, bool last);
    void cancel();                     // wake and stop everything (abort)
    void close();                      // joins the decoder threads
    const char *error() const;
    Stats stats() const;
private:
    Engine *e;
};

} // namespace pinflate
