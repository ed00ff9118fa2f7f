// generate.cu -- synthetic input on the device: pseudo-random data bytes -> convolutional encoder ->
// integer AWGN channel -> quantised soft symbols, one thread per (frame, trellis step).
// Restates the reference's test-input path (src/util.h:8-62 generate_random_bytes/encode_data,
// viterbi/convolutional_encoder_shift_register.h:42-62; channel shaped after
// examples/run_snr_ber.cpp:317-365) with integer-only arithmetic so that this kernel and the host-side
// generator used by the tests produce identical bytes.
#include "kernels.h"

namespace b200v {

namespace {

__device__ __forceinline__ uint64_t hash64(uint64_t seed, uint64_t counter) {
    uint64_t z = seed ^ (counter * 0x9E3779B97F4A7C15ull);
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ unsigned data_byte(uint64_t seed, uint64_t frame, unsigned idx) {
    return (unsigned)(hash64(seed ^ 0xDA7Aull, (frame << 32) + idx) & 0xffu);
}

__global__ void generate_kernel(const GenerateArgs a) {
    const int K = a.c.K, R = a.c.R;
    const unsigned T = a.nbits + (unsigned)(K - 1);
    const size_t total = (size_t)a.n_frames * T;
    for (size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x; gid < total;
         gid += (size_t)gridDim.x * blockDim.x) {
        const unsigned f = (unsigned)(gid / T);
        const unsigned t = (unsigned)(gid - (size_t)f * T);
        const uint64_t frame = a.frame0 + f;

        // shift register after consuming bit t: bit k of reg = data bit (t - k), MSB-first bytes, zero tail
        uint32_t reg = 0;
        {
            const int first = (int)t - (K - 1);           // oldest bit still inside the constraint length
            const int b0 = first < 0 ? 0 : first >> 3, b1 = (int)t >> 3;
            for (int b = b0; b <= b1; b++) {
                const unsigned byte = ((unsigned)b * 8u < a.nbits) ? data_byte(a.seed, frame, (unsigned)b) : 0u;
                if (a.data && (unsigned)b * 8u == t && t < a.nbits)
                    a.data[(size_t)f * a.data_stride + b] = (uint8_t)byte;
                for (int q = 0; q < 8; q++) {
                    const int m = b * 8 + q;              // absolute bit index
                    if (m < first || m > (int)t || m < 0) continue;
                    const unsigned bit = (byte >> (7 - q)) & 1u;
                    reg |= bit << ((int)t - m);
                }
            }
        }

        const uint64_t sym0 = (frame * T + t) * (uint64_t)R;
        unsigned char* row = (unsigned char*)a.syms + (size_t)f * a.sym_stride + (size_t)t * R * a.c.sym_bytes;
        for (int i = 0; i < R; i++) {
            const int p = a.c.poly[i];
            const uint32_t g = (uint32_t)(p < 0 ? -p : p) & ((K >= 32) ? 0xffffffffu : ((1u << K) - 1u));
            const unsigned cbit = (__popc(g & reg) & 1u) ^ (p < 0 ? 1u : 0u);
            long long noise = 0;
            if (a.sigma_q8 != 0) {
                long long gsum = 0;
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const uint64_t h = hash64(a.seed ^ (0xC0FFEEull + (uint64_t)k), sym0 + i);
                    gsum += (long long)(h & 0xffff) + (long long)((h >> 16) & 0xffff) +
                            (long long)((h >> 32) & 0xffff) + (long long)((h >> 48) & 0xffff);
                }
                gsum -= 6ll * 65535ll;
                noise = ((long long)a.sigma_q8 * gsum) >> 16;      // floor division by 65536
            }
            const long long x = (cbit ? a.amp_q8 : -a.amp_q8) + noise;
            long long s = (x + a.mean_q8 + 128) >> 8;              // floor((x + mean + 0.5))
            s = s < a.lo ? a.lo : (s > a.hi ? a.hi : s);
            if (a.c.sym_bytes == 2) ((int16_t*)row)[i] = (int16_t)s;
            else row[i] = (unsigned char)(int)s;                   // u8, or int8 two's complement
        }
    }
}

}  // namespace

cudaError_t launch_generate(const GenerateArgs& a, cudaStream_t st) {
    if (a.n_frames <= 0) return cudaSuccess;
    const size_t total = (size_t)a.n_frames * (a.nbits + (unsigned)(a.c.K - 1));
    size_t blocks = (total + 255) / 256;
    if (blocks > 148u * 64u) blocks = 148u * 64u;
    generate_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace b200v
