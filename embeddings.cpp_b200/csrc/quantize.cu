// Model quantiser on the GPU: f32 / f16 legacy-ggml model file -> q4_0 / q4_1 file, byte-identical to what
// the reference's `quantize` tool writes (models/quantize.cpp:27-268: every 2-D tensor whose name ends in
// "weight" -- the three embedding tables included -- is quantised row by row, everything else is copied;
// block arithmetic ggml.c:942-977 (q4_0) and ggml.c:983-1018 (q4_1), row loop ggml.c:19184-19228).
// lets real checkpoints be quantised on the box that serves them.
//
// One warp per block of 32 weights, one lane per weight.  The arithmetic is IEEE f32 with the roundings the
// reference's compiled code has: d = max / -8 (resp. (max - min) / 15) and id = 1 / d are correctly-rounded
// divisions, `x * id + 8.5f` (resp. `(x - min) * id + 0.5f`) is ONE fused multiply-add (gcc contracts it
// under -O3 -mfma, which is how the reference is built), the (int8_t) cast truncates, and d / m are stored
// as round-to-nearest-even fp16.  The first of equal maxima wins (strict `<` in ggml.c:950-956).
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/bert_b200.h"

namespace {

constexpr int QK = 32;

template <bool F16IN>
__device__ __forceinline__ float load_w(const void * src, size_t i) {
    if (F16IN) return __half2float(reinterpret_cast<const __half *>(src)[i]);
    return reinterpret_cast<const float *>(src)[i];
}

template <bool F16IN, int QTYPE>     // QTYPE 2: q4_0 (18 B / block), 3: q4_1 (20 B / block)
__global__ void quantize_q4_kernel(const void * __restrict__ src, uint8_t * __restrict__ dst, size_t n_blocks) {
    const size_t blk = (size_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (blk >= n_blocks) return;
    const int lane = threadIdx.x & 31;
    const float x = load_w<F16IN>(src, blk * QK + lane);
    uint32_t q;
    constexpr int kBytes = QTYPE == 2 ? 18 : 20;
    uint8_t * out = dst + blk * kBytes;
    if (QTYPE == 2) {
        float am = fabsf(x), mv = x;
        int mi = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float am2 = __shfl_xor_sync(0xffffffffu, am, o), mv2 = __shfl_xor_sync(0xffffffffu, mv, o);
            const int mi2 = __shfl_xor_sync(0xffffffffu, mi, o);
            if (am2 > am || (am2 == am && mi2 < mi)) { am = am2; mv = mv2; mi = mi2; }
        }
        if (am == 0.0f) mv = 0.0f;                       // the reference's `max` keeps its initial +0
        const float d = __fdiv_rn(mv, -8.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        const int v = (int) (int8_t) (int) __fmaf_rn(x, id, 8.5f);
        q = (uint32_t) (v < 15 ? v : 15);
        if (lane == 0) *reinterpret_cast<unsigned short *>(out) = __half_as_ushort(__float2half_rn(d));   // 18 * blk is 2-aligned
    } else {
        float mn = x, mx = x;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        const float d = __fdiv_rn(__fsub_rn(mx, mn), 15.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        const int v = (int) (int8_t) (int) __fmaf_rn(__fsub_rn(x, mn), id, 0.5f);
        q = (uint32_t) (v < 15 ? v : 15);
        if (lane == 0) {
            *reinterpret_cast<unsigned short *>(out) = __half_as_ushort(__float2half_rn(d));
            *reinterpret_cast<unsigned short *>(out + 2) = __half_as_ushort(__float2half_rn(mn));
        }
    }
    // byte j = q[j] | q[j + 16] << 4   (ggml.c:968-974)
    const uint32_t hi = __shfl_down_sync(0xffffffffu, q, 16);
    if (lane < 16) out[(QTYPE == 2 ? 2 : 4) + lane] = (uint8_t) (q | (hi << 4));
}

struct File {
    FILE * f = nullptr;
    ~File() { if (f) fclose(f); }
};

bool rd(FILE * f, void * p, size_t n) { return fread(p, 1, n, f) == n; }
bool wr(FILE * f, const void * p, size_t n) { return fwrite(p, 1, n, f) == n; }

bool ends_with_weight(const std::string & s) {           // the reference's regex ".*weight" (full match)
    return s.size() >= 6 && s.compare(s.size() - 6, 6, "weight") == 0;
}

}  // namespace

extern "C" int32_t bertx_quantize_file(const char * fname_inp, const char * fname_out, int32_t itype) {
    const char * fn = "bertx_quantize_file";
    if (itype != 2 && itype != 3) { fprintf(stderr, "%s: invalid quantization type %d\n", fn, itype); return 1; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        fprintf(stderr, "%s: no CUDA device (the quantiser is a GPU kernel; there is no CPU path)\n", fn);
        return 2;
    }
    File in, out;
    in.f = fopen(fname_inp, "rb");
    if (!in.f) { fprintf(stderr, "%s: failed to open '%s' for reading\n", fn, fname_inp); return 3; }
    out.f = fopen(fname_out, "wb");
    if (!out.f) { fprintf(stderr, "%s: failed to open '%s' for writing\n", fn, fname_out); return 3; }
    uint32_t magic = 0;
    if (!rd(in.f, &magic, 4) || magic != 0x67676d6c) {
        fprintf(stderr, "%s: invalid model file '%s' (bad magic)\n", fn, fname_inp);
        return 4;
    }
    int32_t hp[7];
    if (!rd(in.f, hp, sizeof(hp))) { fprintf(stderr, "%s: truncated header\n", fn); return 4; }
    const int32_t n_vocab = hp[0];
    hp[6] = itype;                                               // quantize.cpp:88
    if (!wr(out.f, &magic, 4) || !wr(out.f, hp, sizeof(hp))) return 5;
    std::vector<char> word;
    for (int i = 0; i < n_vocab; ++i) {
        uint32_t len = 0;
        if (!rd(in.f, &len, 4) || len > (1u << 20)) { fprintf(stderr, "%s: truncated vocab\n", fn); return 4; }
        word.resize(len);
        if (len && !rd(in.f, word.data(), len)) { fprintf(stderr, "%s: truncated vocab\n", fn); return 4; }
        if (!wr(out.f, &len, 4) || (len && !wr(out.f, word.data(), len))) return 5;
    }
    void * d_src = nullptr;
    uint8_t * d_dst = nullptr;
    size_t cap_src = 0, cap_dst = 0;
    std::vector<uint8_t> host, packed;
    int32_t rc = 0;
    size_t total_org = 0, total_new = 0;
    while (true) {
        int32_t n_dims = 0, length = 0, ftype = 0;
        if (!rd(in.f, &n_dims, 4)) break;                        // clean end of file
        if (!rd(in.f, &length, 4) || !rd(in.f, &ftype, 4) || n_dims < 1 || n_dims > 2 || length < 0 || length > 4096) {
            fprintf(stderr, "%s: corrupt tensor header\n", fn); rc = 4; break;
        }
        int32_t ne[2] = {1, 1};
        size_t nel = 1;
        for (int i = 0; i < n_dims; ++i) { if (!rd(in.f, &ne[i], 4) || ne[i] <= 0) { rc = 4; break; } nel *= (size_t) ne[i]; }
        if (rc) { fprintf(stderr, "%s: corrupt tensor shape\n", fn); break; }
        std::string name((size_t) length, '\0');
        if (length && !rd(in.f, &name[0], (size_t) length)) { rc = 4; break; }
        const bool quantize = ends_with_weight(name) && n_dims == 2;
        if (ftype != 0 && ftype != 1) {
            fprintf(stderr, "%s: unsupported ftype %d for %s\n", fn, ftype, quantize ? "integer quantization" : "copy");
            rc = 6; break;
        }
        const size_t bpe = ftype == 0 ? 4 : 2;
        host.resize(nel * bpe);
        if (!rd(in.f, host.data(), host.size())) { fprintf(stderr, "%s: truncated tensor '%s'\n", fn, name.c_str()); rc = 4; break; }
        const int32_t ftype_out = quantize ? itype : ftype;
        if (!wr(out.f, &n_dims, 4) || !wr(out.f, &length, 4) || !wr(out.f, &ftype_out, 4)) { rc = 5; break; }
        for (int i = 0; i < n_dims; ++i) if (!wr(out.f, &ne[i], 4)) rc = 5;
        if (length && !wr(out.f, name.data(), (size_t) length)) rc = 5;
        if (rc) break;
        total_org += host.size();
        if (!quantize) {
            if (!wr(out.f, host.data(), host.size())) { rc = 5; break; }
            total_new += host.size();
            continue;
        }
        if (ne[0] % QK != 0) { fprintf(stderr, "%s: row length %d of '%s' is not a multiple of 32\n", fn, ne[0], name.c_str()); rc = 6; break; }
        const size_t n_blocks = nel / QK, out_bytes = n_blocks * (itype == 2 ? 18 : 20);
        if (host.size() > cap_src) { cudaFree(d_src); cap_src = host.size(); if (cudaMalloc(&d_src, cap_src) != cudaSuccess) { rc = 7; break; } }
        if (out_bytes > cap_dst) { cudaFree(d_dst); cap_dst = out_bytes; if (cudaMalloc((void **) &d_dst, cap_dst) != cudaSuccess) { rc = 7; break; } }
        if (cudaMemcpy(d_src, host.data(), host.size(), cudaMemcpyHostToDevice) != cudaSuccess) { rc = 7; break; }
        const int wpb = 8;                                       // warps per CTA
        const unsigned grid = (unsigned) ((n_blocks + wpb - 1) / wpb);
        if (ftype == 1 && itype == 2) quantize_q4_kernel<true, 2><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        else if (ftype == 1) quantize_q4_kernel<true, 3><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        else if (itype == 2) quantize_q4_kernel<false, 2><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        else quantize_q4_kernel<false, 3><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        packed.resize(out_bytes);
        cudaError_t ce = cudaGetLastError();                     // launch errors are not sticky: ask now
        if (ce == cudaSuccess) ce = cudaMemcpy(packed.data(), d_dst, out_bytes, cudaMemcpyDeviceToHost);
        if (ce != cudaSuccess) {
            fprintf(stderr, "%s: CUDA failure on '%s': %s\n", fn, name.c_str(), cudaGetErrorString(ce));
            rc = 8; break;
        }
        if (!wr(out.f, packed.data(), out_bytes)) { rc = 5; break; }
        total_new += out_bytes;
    }
    cudaFree(d_src);
    cudaFree(d_dst);
    if (rc == 5) fprintf(stderr, "%s: write to '%s' failed\n", fn, fname_out);
    if (rc == 7) fprintf(stderr, "%s: CUDA allocation / copy failed\n", fn);
    if (rc == 0) {
        printf("%s: model size  = %8.2f MB\n", fn, total_org / 1024.0 / 1024.0);
        printf("%s: quant size  = %8.2f MB\n", fn, total_new / 1024.0 / 1024.0);
        fflush(stdout);
    }
    return rc;
}

// ---------------------------------------------------------------------------------------------------------------
// bertx_quantize_rows_noalloc: the same block quantiser for callers that cannot allocate device memory (no counterpart
// in the reference).  The device copies of the source values and of the
// q4 blocks -- what bertx_quantize_file above holds in cudaMalloc'ed buffers -- live in ONE file-scope __device__ array
// of a fixed size, which every GPU that loads this module keeps until the process exits.
//
// Rules that follow from that:
//   * in host code `g_noalloc_scratch` is only a handle for the symbol, not a device address; the address is fetched once
//     per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernel receives;
//   * every call uses the same array and finds the previous call's bytes in it.  Nothing has to start at zero here: the
//     kernel writes every byte of the output region that is copied back, and reads only source bytes this call uploaded;
//   * two calls that overlap on one GPU would race on the array.  Calls of one process are serialised by a host mutex and
//     the entry point returns only after its device work has finished, so that cannot happen through this entry point;
//     anything else that were to borrow the array would have to queue behind it.
namespace {

constexpr size_t kNoAllocMaxElems = (size_t) 1 << 20;                    // values per call (32 768 blocks)
constexpr size_t kNoAllocSrcBytes = kNoAllocMaxElems * 4;                // f32 source
constexpr size_t kNoAllocDstBytes = kNoAllocMaxElems / QK * 20;          // q4_1 blocks (the larger of the two)
constexpr size_t kNoAllocBytes = kNoAllocSrcBytes + kNoAllocDstBytes;    // 4.625 MiB per GPU
static_assert(kNoAllocSrcBytes % 256 == 0, "the block region starts 256-byte aligned");

static __device__ __align__(256) uint8_t g_noalloc_scratch[kNoAllocBytes];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
uint8_t * g_noalloc_addr[kNoAllocMaxDevices] = {};                       // device address of the array, per device ordinal

// Device address of the scratch array on the CURRENT device (looked up once per device).  Caller holds g_noalloc_mu.
uint8_t * noalloc_scratch_addr(const char * fn) {
    int dev = -1;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices) {
        fprintf(stderr, "%s: no usable current CUDA device (%s)\n", fn, ce != cudaSuccess ? cudaGetErrorString(ce) : "ordinal out of range");
        return nullptr;
    }
    if (!g_noalloc_addr[dev]) {
        void * p = nullptr;
        ce = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (ce != cudaSuccess || !p) {
            fprintf(stderr, "%s: cudaGetSymbolAddress failed on device %d: %s\n", fn, dev, cudaGetErrorString(ce));
            return nullptr;
        }
        g_noalloc_addr[dev] = (uint8_t *) p;
    }
    return g_noalloc_addr[dev];
}

}  // namespace

extern "C" int32_t bertx_quantize_rows_noalloc(const void * src, int32_t src_ftype, int64_t n_elems, int32_t itype, void * dst) {
    const char * fn = "bertx_quantize_rows_noalloc";
    if (itype != 2 && itype != 3) { fprintf(stderr, "%s: invalid quantization type %d\n", fn, itype); return 1; }
    if (src_ftype != 0 && src_ftype != 1) { fprintf(stderr, "%s: source type %d is neither f32 (0) nor f16 (1)\n", fn, src_ftype); return 6; }
    if (!src || !dst || n_elems <= 0 || n_elems % QK != 0) {
        fprintf(stderr, "%s: need non-null buffers and a positive multiple of %d values (got %lld)\n", fn, QK, (long long) n_elems);
        return 6;
    }
    if ((uint64_t) n_elems > kNoAllocMaxElems) {
        fprintf(stderr, "%s: %lld values exceed the fixed scratch of %zu values per call (%zu bytes per GPU); split the rows over several calls\n",
                fn, (long long) n_elems, kNoAllocMaxElems, kNoAllocBytes);
        return 9;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        fprintf(stderr, "%s: no CUDA device (the quantiser is a GPU kernel; there is no CPU path)\n", fn);
        return 2;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    uint8_t * scratch = noalloc_scratch_addr(fn);
    if (!scratch) return 7;
    uint8_t * d_src = scratch;
    uint8_t * d_dst = scratch + kNoAllocSrcBytes;
    const size_t n_blocks = (size_t) n_elems / QK, out_bytes = n_blocks * (itype == 2 ? 18 : 20);
    const size_t in_bytes = (size_t) n_elems * (src_ftype == 0 ? 4 : 2);
    cudaError_t ce = cudaMemcpy(d_src, src, in_bytes, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) {
        const int wpb = 8;
        const unsigned grid = (unsigned) ((n_blocks + wpb - 1) / wpb);
        if (src_ftype == 1 && itype == 2) quantize_q4_kernel<true, 2><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        else if (src_ftype == 1) quantize_q4_kernel<true, 3><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        else if (itype == 2) quantize_q4_kernel<false, 2><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        else quantize_q4_kernel<false, 3><<<grid, 32 * wpb>>>(d_src, d_dst, n_blocks);
        ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaMemcpy(dst, d_dst, out_bytes, cudaMemcpyDeviceToHost);   // also waits for the kernel
    if (ce != cudaSuccess) {
        fprintf(stderr, "%s: CUDA failure: %s\n", fn, cudaGetErrorString(ce));
        return 8;
    }
    return 0;
}
