// Synthetic crash-log corpus generator, device side (bench and test INPUT only): fills a device
// buffer with exactly the bytes pm_synth_corpus would write into a host buffer.  BASELINE configs 3
// and 5 need 17-64 GB of text per GPU: that neither fits the host of the 8-GPU box comfortably nor
// should cross PCIe, and it need not - the text of line i is a pure function of (seed, i)
// (pm_synth_core.h), so one thread per line produces it where it is scanned.
//
// Lines are produced in chunks of kChunkLines: (1) every thread measures its line, (2) an exclusive
// prefix sum places the lines, (3) every thread writes its line if it still ends inside the byte
// budget.  The host reads one number per chunk (its byte total) to know where the next chunk starts
// and when the budget is reached - the same "whole lines while they fit" rule as the host generator.
#include <cuda_runtime.h>

#include <cub/cub.cuh>
#include <string>
#include <vector>

#include "pm_synth_core.h"

using namespace pm_synth;

namespace {

constexpr unsigned long long kChunkLines = 1ull << 23;   // ~1.2 GB of text per round, 96 MiB of scratch

__global__ void __launch_bounds__(256) k_syn_len(Params p, unsigned long long first_line, unsigned long long n, uint32_t *__restrict__ len) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    if (i == n) {   // one extra zero so that the exclusive sum of n + 1 items ends in the chunk's byte total
        len[i] = 0;
        return;
    }
    CountSink o;
    gen_line(p, first_line + i, o);
    len[i] = (uint32_t)o.size() + (p.crlf ? 2u : 1u);
}

__global__ void __launch_bounds__(256) k_syn_write(Params p, unsigned long long first_line, unsigned long long n, const unsigned long long *__restrict__ off,
                                                   unsigned long long base, unsigned long long budget, uint8_t *__restrict__ out,
                                                   unsigned long long *__restrict__ fit) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long a = base + off[i], b = base + off[i + 1];
    if (b > budget) return;
    PtrSink o(out + a);
    gen_line(p, first_line + i, o);
    if (p.crlf) o.ch('\r');
    o.ch('\n');
    // exactly one thread of the chunk wrote the last line that fits (the prefix sums grow): it reports how many
    // lines and bytes of this chunk went out
    if (i + 1 == n || base + off[i + 2] > budget) {
        fit[0] = i + 1;
        fit[1] = off[i + 1];
    }
}

struct DeviceStrings {   // witness strings on the device: one pool + a pointer table
    char *pool = nullptr;
    const char **table = nullptr;
    ~DeviceStrings() {
        cudaFree(pool);
        cudaFree(table);
    }
    cudaError_t upload(const char *const *strs, int n, cudaStream_t s) {
        if (!strs || n <= 0) return cudaSuccess;
        std::string blob;
        std::vector<size_t> at(n, (size_t)-1);
        for (int i = 0; i < n; i++)
            if (strs[i]) {
                at[i] = blob.size();
                blob.append(strs[i]);
                blob.push_back('\0');
            }
        if (blob.empty()) blob.push_back('\0');
        cudaError_t e = cudaMalloc(&pool, blob.size());
        if (e != cudaSuccess) return e;
        e = cudaMalloc(&table, sizeof(char *) * n);
        if (e != cudaSuccess) return e;
        std::vector<const char *> ptrs(n);
        for (int i = 0; i < n; i++) ptrs[i] = at[i] == (size_t)-1 ? nullptr : pool + at[i];
        e = cudaMemcpyAsync(pool, blob.data(), blob.size(), cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) return e;
        e = cudaMemcpyAsync(table, ptrs.data(), sizeof(char *) * n, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) return e;
        return cudaStreamSynchronize(s);   // the host vectors go out of scope
    }
};

}  // namespace

// Same contract as pm_synth_corpus (pm_synth.cc) with a DEVICE output buffer of at least target_bytes
// bytes on the current device: returns the bytes written (whole lines, <= target_bytes) or a negative
// cudaError_t, *n_lines_out receives the line count.  Synchronous.
extern "C" int64_t pm_synth_corpus_device_near(uint64_t seed, uint64_t first_line, int64_t target_bytes, double p_hit, const char *const *witnesses,
                                               const char *const *sec_witnesses, int n_wit, int crlf, double p_near, const char *const *near_strings, int n_near,
                                               uint8_t *d_out, int64_t *n_lines_out);

extern "C" int64_t pm_synth_corpus_device(uint64_t seed, uint64_t first_line, int64_t target_bytes, double p_hit, const char *const *witnesses,
                                          const char *const *sec_witnesses, int n_wit, int crlf, uint8_t *d_out, int64_t *n_lines_out) {
    return pm_synth_corpus_device_near(seed, first_line, target_bytes, p_hit, witnesses, sec_witnesses, n_wit, crlf, 0.0, nullptr, 0, d_out, n_lines_out);
}

// The same with near-miss lines (pm_synth_core.h, Params::p_near).
extern "C" int64_t pm_synth_corpus_device_near(uint64_t seed, uint64_t first_line, int64_t target_bytes, double p_hit, const char *const *witnesses,
                                               const char *const *sec_witnesses, int n_wit, int crlf, double p_near, const char *const *near_strings, int n_near,
                                               uint8_t *d_out, int64_t *n_lines_out) {
    if (!d_out || target_bytes < 0) return -1;
    cudaStream_t s = nullptr;
    DeviceStrings wit, swit, near_dev;
    cudaError_t e;
#define SYN_CU(call)                                  \
    do {                                              \
        e = (call);                                   \
        if (e != cudaSuccess) {                       \
            cudaFree(d_len);                          \
            cudaFree(d_off);                          \
            cudaFree(d_tmp);                          \
            cudaFree(d_fit);                          \
            return -(int64_t)e;                       \
        }                                             \
    } while (0)
    uint32_t *d_len = nullptr;
    unsigned long long *d_off = nullptr, *d_fit = nullptr;
    void *d_tmp = nullptr;
    SYN_CU(wit.upload(witnesses, n_wit, s));
    SYN_CU(swit.upload(sec_witnesses, sec_witnesses ? n_wit : 0, s));
    SYN_CU(near_dev.upload(near_strings, n_near, s));
    Params p{seed, p_hit, wit.table, swit.table, n_wit, crlf != 0};
    if (n_near > 0 && near_strings) p.p_near = p_near, p.near = near_dev.table, p.n_near = n_near;
    const unsigned long long budget = (unsigned long long)target_bytes;
    // a line has at least 10 bytes: never measure more lines than the budget can hold
    const unsigned long long chunk = std::min<unsigned long long>(kChunkLines, budget / 10 + 16);
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (const uint32_t *)nullptr, (unsigned long long *)nullptr, (long long)(chunk + 1));
    SYN_CU(cudaMalloc(&d_len, (chunk + 1) * 4));
    SYN_CU(cudaMalloc(&d_off, (chunk + 1) * 8));
    SYN_CU(cudaMalloc(&d_tmp, tmp_bytes));
    SYN_CU(cudaMalloc(&d_fit, 16));
    unsigned long long base = 0, lines = 0, next = first_line;
    for (;;) {
        const unsigned blocks = (unsigned)((chunk + 1 + 255) / 256);
        k_syn_len<<<blocks, 256, 0, s>>>(p, next, chunk, d_len);
        SYN_CU(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_len, d_off, (long long)(chunk + 1), s));
        SYN_CU(cudaMemsetAsync(d_fit, 0, 16, s));
        k_syn_write<<<blocks, 256, 0, s>>>(p, next, chunk, d_off, base, budget, d_out, d_fit);
        unsigned long long fit[2] = {0, 0};
        SYN_CU(cudaMemcpyAsync(fit, d_fit, 16, cudaMemcpyDeviceToHost, s));
        SYN_CU(cudaStreamSynchronize(s));
        SYN_CU(cudaGetLastError());
        base += fit[1];
        lines += fit[0];
        if (fit[0] < chunk) break;   // the budget was reached inside this chunk
        next += chunk;
    }
    cudaFree(d_len);
    cudaFree(d_off);
    cudaFree(d_tmp);
    cudaFree(d_fit);
#undef SYN_CU
    if (n_lines_out) *n_lines_out = (int64_t)lines;
    return (int64_t)base;
}
