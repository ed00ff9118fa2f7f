// depuncture.cu -- depuncturing front-end: expands a punctured symbol stream to the full-rate stream
// the ACS kernels consume. Takes the role of the symbol loop of decode_punctured_symbols
// (williamyang_viterbi_lib/examples/helpers/puncture_code_helpers.h:17-55): output symbol k takes the next
// transmitted symbol where puncture_code[k mod L] is set and `unpunctured_symbol_value` elsewhere; the code index
// starts at 0 for every call (the reference calls it once per segment of a frame, run_punctured_decoder.cpp:260-282).
// One thread per output symbol: rank(k) = (k / L) * ones + prefix[k mod L] transmitted symbols precede position k.
// Pure data movement: HBM-bound (reads `ones/L` of what it writes).
#include "kernels.h"

namespace b200v {

namespace {

template <typename T>
__global__ void depuncture_kernel(const DepunctureArgs a) {
    // frames on grid.x (2^31 - 1 blocks), the symbol blocks of a frame on grid.y (<= 64): gridDim.y is capped at 65535,
    // and the headline batch has 1 Mi frames
    const int frame = blockIdx.x;
    const T* in = reinterpret_cast<const T*>((const unsigned char*)a.in + (size_t)frame * a.in_stride) + a.in_offset;
    T* out = reinterpret_cast<T*>((unsigned char*)a.out + (size_t)frame * a.out_stride) + a.out_offset;
    for (int k = blockIdx.y * blockDim.x + threadIdx.x; k < a.n_out; k += gridDim.y * blockDim.x) {
        const int q = k / a.code_len, j = k - q * a.code_len;
        const bool kept = a.prefix[j + 1] != a.prefix[j];
        out[k] = kept ? in[(size_t)q * a.prefix[a.code_len] + a.prefix[j]] : (T)a.fill;
    }
}

// the same expansion through a precomputed gather table (b200v_batch_set_puncturing): out[k] = in[idx[k]] or fill.
// Used by the kernel families whose symbol fetch does not take the table itself.
template <typename T>
__global__ void gather_kernel(const unsigned char* in, size_t in_stride, unsigned char* out, size_t out_stride,
                              const int* idx, int n_out, int fill, int frame0) {
    const int frame = frame0 + blockIdx.x;
    const T* pi = reinterpret_cast<const T*>(in + (size_t)frame * in_stride);
    T* po = reinterpret_cast<T*>(out + (size_t)frame * out_stride);
    for (int k = blockIdx.y * blockDim.x + threadIdx.x; k < n_out; k += gridDim.y * blockDim.x) {
        const int ix = __ldg(idx + k);
        po[k] = ix >= 0 ? pi[ix] : (T)fill;
    }
}

}  // namespace

cudaError_t launch_depuncture_gather(const void* in, size_t in_stride, void* out, size_t out_stride, const int* idx,
                                     int n_out, int fill, int sym_bytes, int frame0, int count, cudaStream_t st) {
    if (count <= 0 || n_out <= 0) return cudaSuccess;
    int bx = (n_out + 255) / 256;
    if (bx > 64) bx = 64;
    const dim3 grid((unsigned)count, (unsigned)bx);
    if (sym_bytes == 2) gather_kernel<int16_t><<<grid, 256, 0, st>>>((const unsigned char*)in, in_stride, (unsigned char*)out, out_stride, idx, n_out, fill, frame0);
    else gather_kernel<uint8_t><<<grid, 256, 0, st>>>((const unsigned char*)in, in_stride, (unsigned char*)out, out_stride, idx, n_out, fill, frame0);
    return cudaGetLastError();
}

cudaError_t launch_depuncture(const DepunctureArgs& a, int sym_bytes, int n_frames, cudaStream_t st) {
    if (n_frames <= 0 || a.n_out <= 0) return cudaSuccess;
    int bx = (a.n_out + 255) / 256;
    if (bx > 64) bx = 64;
    const dim3 grid((unsigned)n_frames, (unsigned)bx);
    if (sym_bytes == 2) depuncture_kernel<int16_t><<<grid, 256, 0, st>>>(a);
    else depuncture_kernel<uint8_t><<<grid, 256, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace b200v
