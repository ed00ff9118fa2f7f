// ubench_int.cu -- calibrates the integer-pipe roofline on the box: sustained per-SM issue rate
// of the SASS instructions the ACS kernels are built from. Each kernel runs ILP independent dependency
// chains per thread so that latency is hidden; result = warp-instructions * 32 lanes / clk / SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_int tools/ubench_int.cu && ./ubench_int
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

constexpr int ILP = 8;
constexpr int ITERS = 4096;

#define DEFINE_KERNEL(NAME, ...)                                                      \
    __global__ void NAME(unsigned* out, unsigned seed) {                              \
        unsigned a[ILP], b = seed * 3u + 1u, c = seed ^ 0x00ff00ffu;                  \
        for (int i = 0; i < ILP; i++) a[i] = threadIdx.x * 2654435761u + i + seed;    \
        for (int it = 0; it < ITERS; it++) {                                          \
            _Pragma("unroll") for (int i = 0; i < ILP; i++) { __VA_ARGS__; }               \
        }                                                                             \
        unsigned r = 0;                                                               \
        for (int i = 0; i < ILP; i++) r ^= a[i];                                      \
        r ^= b;                                                                       \
        if (r == 0x12345678u) out[0] = r;                                             \
    }

#define Y a[(i + 3) & (ILP - 1)]
#define H2(x) (*reinterpret_cast<__half2*>(&(x)))
DEFINE_KERNEL(k_iadd, a[i] = a[i] + Y)
DEFINE_KERNEL(k_lop3, a[i] = (a[i] & Y) ^ c)
DEFINE_KERNEL(k_shf, a[i] = __funnelshift_r(a[i], Y, 1))
DEFINE_KERNEL(k_prmt, a[i] = __byte_perm(a[i], Y, 0x3120))
DEFINE_KERNEL(k_imad, a[i] = a[i] * b + Y)
DEFINE_KERNEL(k_viaddmin_u16x2, a[i] = __viaddmin_u16x2(a[i], Y, c))
DEFINE_KERNEL(k_viaddmin_s16x2, a[i] = __viaddmin_s16x2(a[i], Y, c))
DEFINE_KERNEL(k_vimin3_u16x2, a[i] = __vimin3_u16x2(a[i], Y, c))
DEFINE_KERNEL(k_vadd2, a[i] = __vadd2(a[i], Y))
DEFINE_KERNEL(k_vabsdiff4, a[i] = __vabsdiffu4(a[i], Y))
// min + add alternating (cannot fold into VIMNMX3): 1 VIMNMX + 1 VIADD.16x2 per body
DEFINE_KERNEL(k_vimin_vadd2, a[i] = __vadd2(__vminu2(a[i], Y), c))
// min with both predicate outputs consumed (decision path): VIMNMX.U16x2 P0,P1 + SEL/LOP per body
DEFINE_KERNEL(k_vibmin_pred, { bool p, q; a[i] = __vibmin_u16x2(a[i], Y, &p, &q) + c; b += (unsigned)p + 2u * (unsigned)q; })
DEFINE_KERNEL(k_hadd2, { __half2 h = __hadd2(H2(a[i]), H2(Y)); a[i] = *reinterpret_cast<unsigned*>(&h); })
DEFINE_KERNEL(k_hmin2, { __half2 h = __hadd2(__hmin2(H2(a[i]), H2(Y)), H2(c)); a[i] = *reinterpret_cast<unsigned*>(&h); })
// mixes: does the FMA pipe (IMAD / HADD2) issue alongside the ALU pipe?
DEFINE_KERNEL(k_mix_imad_lop3, { if (i & 1) a[i] = a[i] * b + Y; else a[i] = (a[i] & Y) ^ c; })
DEFINE_KERNEL(k_mix_imad_viaddmin, { if (i & 1) a[i] = a[i] * b + Y; else a[i] = __viaddmin_u16x2(a[i], Y, c); })
DEFINE_KERNEL(k_mix_hadd2_viaddmin, { if (i & 1) { __half2 h = __hadd2(H2(a[i]), H2(Y)); a[i] = *reinterpret_cast<unsigned*>(&h); } else a[i] = __viaddmin_u16x2(a[i], Y, c); })
DEFINE_KERNEL(k_mix_lop3_viaddmin, { if (i & 1) a[i] = (a[i] & Y) ^ c; else a[i] = __viaddmin_u16x2(a[i], Y, c); })

// ---- round 2: candidates for taking the decision collection off the issue budget -----------------------------------
// HSET2 with a mask result (0xffff / 0 per half): which pipe, and does it issue beside the DPX ops and IMAD?
__device__ __forceinline__ unsigned hne2m(unsigned a, unsigned b) { unsigned d; asm("set.ne.u32.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ unsigned imad(unsigned a, unsigned b, unsigned c) { unsigned d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ unsigned iadd3(unsigned a, unsigned b, unsigned z) { unsigned d; asm("{\n\t.reg .u32 t;\n\tadd.u32 t, %1, %2;\n\tadd.u32 %0, t, %3;\n\t}" : "=r"(d) : "r"(a), "r"(b), "r"(z)); return d; }
DEFINE_KERNEL(k_hset2, a[i] = hne2m(a[i], Y) ^ 0u)
DEFINE_KERNEL(k_mix_hset2_viaddmin, { if (i & 1) a[i] = hne2m(a[i], Y); else a[i] = __viaddmin_u16x2(a[i], Y, c); })
DEFINE_KERNEL(k_mix_hset2_imad, { if (i & 1) a[i] = hne2m(a[i], Y); else a[i] = imad(a[i], b, Y); })
DEFINE_KERNEL(k_mix_hset2_hadd2, { if (i & 1) a[i] = hne2m(a[i], Y); else { __half2 h = __hadd2(H2(a[i]), H2(Y)); a[i] = *reinterpret_cast<unsigned*>(&h); } })
// three pipes at once: DPX (ALU) + IMAD (FMA heavy) + HSET2
DEFINE_KERNEL(k_mix3_viaddmin_imad_hset2, { if ((i & 3) == 0) a[i] = imad(a[i], b, Y); else if ((i & 3) == 2) a[i] = hne2m(a[i], Y); else a[i] = __viaddmin_u16x2(a[i], Y, c); })
// the fused ACS body for one packed pair of new states: t = lo path (add), n = min(hi + Wc, t) (VIADDMNMX), mask = (n != t)
// (HSET2), S = 2 S + mask (Horner on IMAD or IADD3): 4 instructions per 2 new states and 2 decisions
#define ACS_FUSED(ADD_T, HORNER)                                                                  \
    { const unsigned t = ADD_T; const unsigned n = __viaddmin_u16x2(Y, c, t); const unsigned m = hne2m(n, t); \
      b = HORNER; a[i] = n; }
DEFINE_KERNEL(k_acs_fused_fma_t_alu_h, ACS_FUSED(imad(a[i], seed | 1u, c), iadd3(b, b, m)))
DEFINE_KERNEL(k_acs_fused_alu_t_fma_h, ACS_FUSED(iadd3(a[i], c, seed & 0u), imad(b, 2u + (seed & 0u), m)))
DEFINE_KERNEL(k_acs_fused_alu_t_alu_h, ACS_FUSED(iadd3(a[i], c, seed & 0u), iadd3(b, b, m)))
DEFINE_KERNEL(k_acs_fused_fma_t_fma_h, ACS_FUSED(imad(a[i], seed | 1u, c), imad(b, 2u + (seed & 0u), m)))
// today's body: two adds, VIMNMX with predicate outputs, two predicated adds: 5 instructions per 2 new states
DEFINE_KERNEL(k_acs_pred, { const unsigned t = imad(a[i], seed | 1u, c); const unsigned u = iadd3(Y, c, seed & 0u); bool p, q;
                            a[i] = __vibmin_u16x2(t, u, &p, &q); if (p) b += 1u << i; if (q) b += 0x10000u << i; })

// ---- round 2b: is there a third pipe? HFMA2 / FFMA beside IMAD (FMA heavy) and the DPX ops (ALU) -------------------------
__device__ __forceinline__ unsigned hfma2u(unsigned a, unsigned b, unsigned c) { unsigned d; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ unsigned ffmau(unsigned a, unsigned b, unsigned c) { return __float_as_uint(fmaf(__uint_as_float(a), __uint_as_float(b), __uint_as_float(c))); }
__device__ __forceinline__ unsigned hne2h(unsigned a, unsigned b) { unsigned d; asm("set.ne.f16x2.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
DEFINE_KERNEL(k_hfma2, a[i] = hfma2u(a[i], c, Y))
DEFINE_KERNEL(k_ffma, a[i] = ffmau(a[i], c, Y))
DEFINE_KERNEL(k_mix_hfma2_imad, { if (i & 1) a[i] = hfma2u(a[i], c, Y); else a[i] = imad(a[i], b, Y); })
DEFINE_KERNEL(k_mix_ffma_imad, { if (i & 1) a[i] = ffmau(a[i], c, Y); else a[i] = imad(a[i], b, Y); })
DEFINE_KERNEL(k_mix_ffma_hfma2, { if (i & 1) a[i] = ffmau(a[i], c, Y); else a[i] = hfma2u(a[i], c, Y); })
DEFINE_KERNEL(k_mix_ffma_viaddmin, { if (i & 1) a[i] = ffmau(a[i], c, Y); else a[i] = __viaddmin_u16x2(a[i], Y, c); })
DEFINE_KERNEL(k_mix4_2viaddmin_imad_hfma2, { if ((i & 3) == 0) a[i] = imad(a[i], b, Y); else if ((i & 3) == 2) a[i] = hfma2u(a[i], c, Y); else a[i] = __viaddmin_u16x2(a[i], Y, c); })
DEFINE_KERNEL(k_mix4_2viaddmin_imad_ffma, { if ((i & 3) == 0) a[i] = imad(a[i], b, Y); else if ((i & 3) == 2) a[i] = ffmau(a[i], c, Y); else a[i] = __viaddmin_u16x2(a[i], Y, c); })
// fused ACS body with the fold on HFMA2: mask as f16 1.0 / 0.0, h = 2 h + m exact below 2^11
#define ACS_FUSED_H(ADD_T)                                                                  \
    { const unsigned t = ADD_T; const unsigned n = __viaddmin_u16x2(Y, c, t); const unsigned m = hne2h(n, t); \
      b = hfma2u(b, 0x40004000u, m); a[i] = n; }
DEFINE_KERNEL(k_acs_fused_fma_t_hfma2_h, ACS_FUSED_H(imad(a[i], seed | 1u, c)))
DEFINE_KERNEL(k_acs_fused_alu_t_hfma2_h, ACS_FUSED_H(iadd3(a[i], c, seed & 0u)))
DEFINE_KERNEL(k_acs_fused_add_t_hfma2_h, ACS_FUSED_H(a[i] + c))
DEFINE_KERNEL(k_acs_fused_add_t_fma_h, ACS_FUSED(a[i] + c, imad(b, 2u + (seed & 0u), m)))

typedef void (*kern_t)(unsigned*, unsigned);

static void run(const char* name, kern_t k, unsigned* d_out, int sms, double clk_ghz, double ops_per_body) {
    const int threads = 512, blocks = sms * 4;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<blocks, threads>>>(d_out, 1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 5; r++) k<<<blocks, threads>>>(d_out, r + 2);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double lane_ops = 5.0 * blocks * threads * (double)ITERS * ILP * ops_per_body;
    const double per_s = lane_ops / (ms * 1e-3);
    printf("{\"kernel\": \"%s\", \"ms\": %.3f, \"lane_ops_per_s\": %.4e, \"lanes_per_clk_per_sm_at_%.0fMHz\": %.1f}\n",
           name, ms, per_s, clk_ghz * 1e3, per_s / (sms * clk_ghz * 1e9));
}

int main() {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) { printf("no device\n"); return 1; }
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double ghz = clk_khz * 1e-6;
    unsigned* d_out;
    cudaMalloc(&d_out, 4);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_ghz\": %.3f}\n", p.name, p.multiProcessorCount, ghz);
#define RUN(K, OPS) run(#K, K, d_out, p.multiProcessorCount, ghz, OPS)
    RUN(k_iadd, 1); RUN(k_lop3, 1); RUN(k_shf, 1); RUN(k_prmt, 1); RUN(k_imad, 1);
    RUN(k_viaddmin_u16x2, 1); RUN(k_viaddmin_s16x2, 1); RUN(k_vimin3_u16x2, 1); RUN(k_vadd2, 1);
    RUN(k_vabsdiff4, 1); RUN(k_vimin_vadd2, 2); RUN(k_vibmin_pred, 1); RUN(k_hadd2, 1); RUN(k_hmin2, 2);
    RUN(k_mix_imad_lop3, 1); RUN(k_mix_imad_viaddmin, 1); RUN(k_mix_hadd2_viaddmin, 1); RUN(k_mix_lop3_viaddmin, 1);
    RUN(k_hset2, 1); RUN(k_mix_hset2_viaddmin, 1); RUN(k_mix_hset2_imad, 1); RUN(k_mix_hset2_hadd2, 1);
    RUN(k_mix3_viaddmin_imad_hset2, 1);
    RUN(k_acs_fused_fma_t_alu_h, 4); RUN(k_acs_fused_alu_t_fma_h, 4); RUN(k_acs_fused_alu_t_alu_h, 4); RUN(k_acs_fused_fma_t_fma_h, 4);
    RUN(k_acs_pred, 5);
    RUN(k_hfma2, 1); RUN(k_ffma, 1); RUN(k_mix_hfma2_imad, 1); RUN(k_mix_ffma_imad, 1); RUN(k_mix_ffma_hfma2, 1);
    RUN(k_mix_ffma_viaddmin, 1); RUN(k_mix4_2viaddmin_imad_hfma2, 1); RUN(k_mix4_2viaddmin_imad_ffma, 1);
    RUN(k_acs_fused_fma_t_hfma2_h, 4); RUN(k_acs_fused_alu_t_hfma2_h, 4); RUN(k_acs_fused_add_t_hfma2_h, 4);
    RUN(k_acs_fused_add_t_fma_h, 4);
    cudaFree(d_out);
    return 0;
}
