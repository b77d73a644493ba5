// pm_line_index_noalloc: the line table of one log, computed without allocating anything.
//
// Who this is for: code that is handed device buffers but is not allowed to create any --
// a plugin running inside a host that owns every byte of GPU memory through its own pool, a
// callback on a path where cudaMalloc's implicit device synchronisation is unacceptable, or a
// process whose memory budget was fixed at start-up.  Every other entry point of this library
// takes its working memory from the context's pool (which calls cudaMalloc when it has to grow);
// this one does not: its scratch (newline counts per 2 KiB tile and their prefix sums) lives in a
// fixed array that is part of the loaded module.  The price is a size limit and the loss of
// re-entrancy, both described at the entry point below.
//
// The kernels are separate from the batch pipeline's (pm_kernels.cu) on purpose: a caller's
// buffer has no '\n' padding around it and no slack behind it, so nothing here may read outside
// [text, text + n).
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <string>

#include "podmortem_scan.h"

namespace {

constexpr unsigned long long kNaMaxBytes = PM_NOALLOC_MAX_BYTES;   // 64 MiB of log text
constexpr int kNaTile = 2048;
constexpr unsigned kNaMaxTiles = (unsigned)(kNaMaxBytes / kNaTile);                 // 32768
constexpr size_t kNaHeader = 256;
constexpr size_t kNaScratchBytes = kNaHeader + (size_t)kNaMaxTiles * 4 + ((size_t)kNaMaxTiles + 1) * 4 + 252;

// The scratch itself.  It belongs to the module image: every GPU that loads this library holds
// these ~257 KiB until the process exits.
static __device__ __align__(256) unsigned char g_na_scratch[kNaScratchBytes];

struct NaHeader {
    unsigned long long eff_end;   // length without trailing "\r?\n" separators
    unsigned long long n_lines;
};

__device__ __forceinline__ NaHeader *na_header(unsigned char *s) { return reinterpret_cast<NaHeader *>(s); }
__device__ __forceinline__ unsigned *na_tile_nl(unsigned char *s) { return reinterpret_cast<unsigned *>(s + kNaHeader); }
__device__ __forceinline__ unsigned *na_tile_prefix(unsigned char *s) { return reinterpret_cast<unsigned *>(s + kNaHeader + (size_t)kNaMaxTiles * 4); }

// Step 1 (one CTA): strip trailing separators the way Java's split does (SPEC.md section 2) and
// reset the header.  The previous call's header is still in the scratch, so everything that
// must start from a known value is written here, not assumed.
//
// Java drops trailing empty lines, i.e. a suffix of the form ("\r"? "\n")*.  Read backwards that
// is: the effective end sits just behind the LAST byte that is kept, and a byte is kept when it is
// neither '\r' nor '\n', or when it is a '\r' that is not directly followed by '\n' (a lone '\r'
// is text, not a separator).  The CTA searches for that byte from the end, 4 KiB per step, so a
// log that ends in megabytes of blank lines costs microseconds per megabyte instead of one
// dependent byte load per separator.
constexpr int kNaBeginThreads = 1024, kNaBeginPer = 4;

__global__ void __launch_bounds__(kNaBeginThreads) k_na_begin(const unsigned char *__restrict__ text, unsigned long long n, unsigned char *scratch) {
    __shared__ long long warp_best[kNaBeginThreads / 32];
    __shared__ long long found_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr unsigned long long kStep = (unsigned long long)kNaBeginThreads * kNaBeginPer;
    long long found = -1;
    for (unsigned long long hi = n; hi > 0 && found < 0; hi = hi > kStep ? hi - kStep : 0) {
        const unsigned long long lo = hi > kStep ? hi - kStep : 0;
        long long best = -1;
#pragma unroll
        for (int k = 0; k < kNaBeginPer; k++) {
            const unsigned long long i = lo + (unsigned long long)threadIdx.x * kNaBeginPer + k;
            if (i < hi) {
                const unsigned char c = text[i];
                const bool keep = (c != '\r' && c != '\n') || (c == '\r' && (i + 1 == n || text[i + 1] != '\n'));
                if (keep) best = (long long)i;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
        if (lane == 0) warp_best[warp] = best;
        __syncthreads();
        if (warp == 0) {
            long long b2 = warp_best[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) b2 = max(b2, __shfl_xor_sync(0xFFFFFFFFu, b2, o));
            if (lane == 0) found_s = b2;
        }
        __syncthreads();
        found = found_s;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        NaHeader *h = na_header(scratch);
        h->eff_end = (unsigned long long)(found + 1);
        h->n_lines = 0;
    }
}

// Step 2: '\n' count of every 2 KiB tile of text[0, eff_end).  One warp per tile, byte loads
// (the caller's pointer carries no alignment promise).
__global__ void __launch_bounds__(256) k_na_count(const unsigned char *__restrict__ text, unsigned char *scratch) {
    const unsigned long long eff = na_header(scratch)->eff_end;
    const unsigned n_tiles = (unsigned)((eff + kNaTile - 1) / kNaTile);
    const int lane = threadIdx.x & 31;
    unsigned *tile_nl = na_tile_nl(scratch);
    for (unsigned t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < n_tiles; t += gridDim.x * (blockDim.x >> 5)) {
        const unsigned long long base = (unsigned long long)t * kNaTile;
        unsigned c = 0;
        for (int k = lane; k < kNaTile; k += 32)
            if (base + k < eff && text[base + k] == '\n') c++;
        c = __reduce_add_sync(0xFFFFFFFFu, c);
        if (lane == 0) tile_nl[t] = c;
    }
}

// Step 3 (one CTA): exclusive prefix of the tile counts, and the line count.
__global__ void __launch_bounds__(1024) k_na_prefix(unsigned char *scratch, unsigned long long n) {
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry_s;
    NaHeader *h = na_header(scratch);
    const unsigned long long eff = h->eff_end;
    const unsigned n_tiles = (unsigned)((eff + kNaTile - 1) / kNaTile);
    const unsigned *tile_nl = na_tile_nl(scratch);
    unsigned *prefix = na_tile_prefix(scratch);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (unsigned b0 = 0; b0 < n_tiles; b0 += 1024) {
        unsigned i = b0 + threadIdx.x, v = i < n_tiles ? tile_nl[i] : 0, inc = v;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned y = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += y;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            unsigned w = wsum[lane], winc = w;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned y = __shfl_up_sync(0xFFFFFFFFu, winc, o);
                if (lane >= o) winc += y;
            }
            wsum[lane] = winc - w;
        }
        __syncthreads();
        unsigned carry = carry_s;
        if (i < n_tiles) prefix[i] = carry + wsum[warp] + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + wsum[31] + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        // SPEC.md section 2: "" is one empty line; a log made only of separators has no lines
        h->n_lines = n == 0 ? 1 : (eff == 0 ? 0 : (unsigned long long)carry_s + 1);
    }
}

// Step 4: scatter.  Line 0 starts at 0; newline number r ends line r and starts line r + 1.
__global__ void __launch_bounds__(256) k_na_scatter(const unsigned char *__restrict__ text, unsigned char *scratch, long long *__restrict__ out,
                                                    unsigned long long cap, unsigned long long *n_lines_out) {
    const NaHeader *h = na_header(scratch);
    const unsigned long long eff = h->eff_end, n_lines = h->n_lines;
    const unsigned n_tiles = (unsigned)((eff + kNaTile - 1) / kNaTile);
    const unsigned *prefix = na_tile_prefix(scratch);
    const int lane = threadIdx.x & 31;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *n_lines_out = n_lines;
        if (n_lines > 0 && cap > 0) out[0] = 0;
    }
    for (unsigned t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < n_tiles; t += gridDim.x * (blockDim.x >> 5)) {
        const unsigned long long base = (unsigned long long)t * kNaTile;
        unsigned long long rank = prefix[t];
        for (int k0 = 0; k0 < kNaTile; k0 += 32) {
            const unsigned long long pos = base + k0 + lane;
            const bool nl = pos < eff && text[pos] == '\n';
            const unsigned b = __ballot_sync(0xFFFFFFFFu, nl);
            if (nl) {
                unsigned long long r = rank + __popc(b & ((1u << lane) - 1u));
                if (r + 1 < n_lines && r + 1 < cap) out[r + 1] = (long long)(pos + 1);
            }
            rank += __popc(b);
        }
    }
}

thread_local std::string t_err;

int na_fail(int status, const char *fmt, ...) {
    char buf[384];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_err = buf;
    return status;
}

// Device address of the scratch, per device.  The symbol's name is only a host-side stand-in:
// handing `g_na_scratch` itself to a kernel from host code would pass a host address.
// The address is looked up once per device ordinal and kept for the life of the process; a
// cudaDeviceReset() unloads the module it points into, so after a reset this entry point must not be
// used again in the same process (documented at the declaration).
constexpr int kNaMaxDevices = 64;
std::mutex g_na_mu;
unsigned char *g_na_ptr[kNaMaxDevices] = {nullptr};

int na_scratch_for_current_device(unsigned char **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return na_fail(PM_E_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kNaMaxDevices) return na_fail(PM_E_INVALID, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> g(g_na_mu);
    if (!g_na_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_na_scratch);   // resolves against the CURRENT device's copy of the module
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return na_fail(PM_E_CUDA, "cudaGetSymbolAddress(scratch): %s", cudaGetErrorString(e));
        }
        g_na_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_na_ptr[dev];
    return PM_OK;
}

}  // namespace

extern "C" const char *pm_noalloc_last_error(void) { return t_err.c_str(); }

extern "C" int pm_line_index_noalloc(const uint8_t *d_text, uint64_t n_bytes, int64_t *d_line_start, uint64_t cap, uint64_t *d_n_lines, void *stream) {
    if (n_bytes > kNaMaxBytes)
        return na_fail(PM_E_INVALID, "pm_line_index_noalloc: %llu bytes exceed the fixed scratch limit of %llu bytes (use pm_batch_line_index)",
                       (unsigned long long)n_bytes, (unsigned long long)kNaMaxBytes);
    if ((!d_text && n_bytes) || !d_n_lines || (!d_line_start && cap)) return na_fail(PM_E_INVALID, "pm_line_index_noalloc: null argument");
    unsigned char *scratch = nullptr;
    int rc = na_scratch_for_current_device(&scratch);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // An error left behind by some earlier, unrelated call on this thread must neither be reported as a
    // failure of these launches nor be cleared by them: look without clearing, and refuse to launch on top of it.
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess)
        return na_fail(PM_E_CUDA, "pm_line_index_noalloc: an earlier CUDA error is pending on this thread (%s); nothing was launched", cudaGetErrorString(e));
    const unsigned n_tiles = (unsigned)((n_bytes + kNaTile - 1) / kNaTile);
    const unsigned blocks = n_tiles ? (n_tiles + 7) / 8 : 1;
    k_na_begin<<<1, kNaBeginThreads, 0, s>>>(d_text, n_bytes, scratch);
    k_na_count<<<blocks < 1184 ? blocks : 1184, 256, 0, s>>>(d_text, scratch);
    k_na_prefix<<<1, 1024, 0, s>>>(scratch, n_bytes);
    k_na_scatter<<<blocks < 1184 ? blocks : 1184, 256, 0, s>>>(d_text, scratch, reinterpret_cast<long long *>(d_line_start), cap,
                                                                 reinterpret_cast<unsigned long long *>(d_n_lines));
    e = cudaGetLastError();   // no error was pending before the four launches above: this one is theirs
    if (e != cudaSuccess) return na_fail(PM_E_CUDA, "pm_line_index_noalloc launch: %s", cudaGetErrorString(e));
    return PM_OK;
}
