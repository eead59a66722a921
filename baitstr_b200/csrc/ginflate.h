// ginflate.h -- gzip inflate on the GPU, host interface (ginflate.cu; device code in ginflate_dev.cuh).
//
// A "*.gz" file is processed in WINDOWS of compressed bytes.  A window goes through stages; the
// stages marked (ordered) must run in window order for one file, the others in any order and on
// any device -- the state that links the windows of a file (FileState: where the next run must
// start, the last 32 KiB of text, the CRC of the open gzip member) lives in host memory, so the
// windows of ONE file can be dealt to SEVERAL GPUs:
//   decode    upload the window (+ look-ahead) and decode all ranges from guessed block starts
//   chain     (ordered) accept the runs that continue the stream, re-decode what does not fit,
//             propagate the 32 KiB history through the runs
//   crc       CRC-32 of the member pieces on the device; crc_fold (ordered) checks CRC-32 and ISIZE
//   resolve   all symbols to bytes, into a device buffer of the caller (the text never visits the host)
// Replaces gzip.open of /root/reference/src/fastq.pyx:50-51 for the select stage's inputs.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace ginflate {

struct FileState {
    uint64_t next_bit = 0;          // bit position in the FILE where the next window's first run starts
    uint8_t hist[32768];            // the last 32 KiB of inflated text
    uint32_t crc = 0;               // CRC-32 of the open gzip member so far
    uint64_t len = 0;               // its length so far
    uint64_t members = 0, out_bytes = 0, redone = 0, ranges = 0;
    bool ended = false;             // the last member ended at the end of the file
    FileState() { for (int i = 0; i < 32768; i++) hist[i] = 0; }
};

// Geometry: the block-start search costs about the same for every range (it ends at the first
// block, ~half a block in), decoding grows with the range, and history propagation takes one step
// per range -- so ranges are large, and a window holds enough of them to fill the GPU's ~3000 warp
// slots (profiles/r02_ginflate_stages.txt).
constexpr size_t DEFAULT_WINDOW = (size_t)256 << 20, DEFAULT_RANGE = (size_t)128 << 10, LOOK_AHEAD = (size_t)4 << 20;

struct DeviceImpl;

class Device {
public:
    // working set for windows of up to window_bytes (+ look_ahead) compressed bytes, cut into
    // ranges of range_bytes.  The CUDA device must be current.  nullptr + message on failure.
    static Device *create(size_t window_bytes, size_t look_ahead, size_t range_bytes, char *err, size_t errlen);
    ~Device();
    // comp[0, nbytes) = file bytes [file_off, file_off + nbytes): the window's own_bytes first, then
    // look-ahead.  at_eof: the buffer ends where the file ends.
    int decode(const uint8_t *comp, size_t nbytes, size_t own_bytes, uint64_t file_off, bool at_eof, cudaStream_t st);
    int chain(FileState &fs, cudaStream_t st);              // (ordered)
    int crc_launch(cudaStream_t st);
    int crc_fold(FileState &fs, cudaStream_t st);           // (ordered)
    uint64_t text_bytes() const;                            // after chain()
    int resolve(uint8_t *d_text, cudaStream_t st);          // enqueue only
    const char *error() const;
    size_t device_bytes() const;
    void kernel_ms(double *upload, double *decode, double *propagate) const;   // CUDA-event times summed over the windows so far
private:
    DeviceImpl *p = nullptr;
};

}  // namespace ginflate
