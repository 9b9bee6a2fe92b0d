// Developer check (GPU box): issue cost per SMSP of single instructions the epilogues use or could use, as cycles per
// warp-instruction with 4 warps per SMSP and 8 independent dependency chains per thread (so latency is hidden and the
// figure is the pipe's issue interval).  Pure instruction loops (asm volatile, no helper ops).
#include <cstdio>
#include <cstdint>
typedef unsigned long long u64;
template <int OP>
__global__ void k(int iters, long long *cyc, uint32_t *sink, uint32_t seed, float fb, float fc) {
  uint32_t a[8], b[8];
  u64 w[8];
  __shared__ uint32_t sm[1024];
  sm[threadIdx.x & 1023] = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    a[i] = (seed + threadIdx.x * 8 + i) & 0xFFFFFu;
    w[i] = ((u64)a[i] << 32) | a[i];
    b[i] = a[i] + 3;
  }
  u64 fb2, fc2;
  asm volatile("mov.b64 %0, {%1, %1};" : "=l"(fb2) : "f"(fb));
  asm volatile("mov.b64 %0, {%1, %1};" : "=l"(fc2) : "f"(fc));
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(a[i]) : "f"(fb), "f"(fc));                  // FFMA on denormal-range patterns
      if (OP == 1) asm volatile("fma.rn.sat.f32 %0, %0, %1, %2;" : "+r"(a[i]) : "f"(fb), "f"(fc));              // FFMA.SAT
      if (OP == 2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(w[i]) : "l"(fb2), "l"(fc2));              // FFMA2
      if (OP == 3) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(seed), "r"(seed));              // IMAD
      if (OP == 4) asm volatile("max.s32 %0, %0, %1;" : "+r"(a[i]) : "r"(seed));                                 // VIMNMX
      if (OP == 5) asm volatile("shr.u32 %0, %1, %0;" : "+r"(a[i]) : "r"(seed));                                 // SHF
      if (OP == 6) asm volatile("add.s32 %0, %0, %1;" : "+r"(a[i]) : "r"(seed));                                 // IADD3
      if (OP == 7) asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(seed), "r"(seed));             // IDP.4A
      if (OP == 8) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(seed), "r"(a[(i + 1) & 7]));       // PRMT, selector from a register
      if (OP == 9) asm volatile("max.f32 %0, %0, %1;" : "+r"(a[i]) : "f"(fb));                                   // FMNMX
      if (OP == 10) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a[i]) : "r"((uint32_t)__cvta_generic_to_shared(sm) + ((a[i] & 1023u) << 2)));   // LDS (dependent address)
      if (OP == 11) asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(seed));                              // VIMNMX.U16x2
      if (OP == 12) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(seed), "r"(seed)); // I2IP
      if (OP == 13) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(seed), "r"(seed));          // LOP3
      if (OP == 14) {                                                                                            // FFMA2 + SHF side by side (separate pipes?)
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(w[i]) : "l"(fb2), "l"(fc2));
        asm volatile("shr.u32 %0, %1, %0;" : "+r"(a[i]) : "r"(seed));
      }
      if (OP == 15) {                                                                                            // FFMA2 + IMAD side by side (same pipe?)
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(w[i]) : "l"(fb2), "l"(fc2));
        asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(seed), "r"(seed));
      }
      if (OP == 16) {                                                                                            // SHF + IMAD side by side
        asm volatile("shr.u32 %0, %1, %0;" : "+r"(a[i]) : "r"(seed));
        asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(b[i]) : "r"(seed), "r"(seed));
      }
      if (OP == 17) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(w[i]) : "l"(fb2));                           // FMUL2
      if (OP == 18) asm volatile("add.rz.f32 %0, %0, %1;" : "+r"(a[i]) : "f"(fc));                               // FADD.RZ
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= a[i] ^ b[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32);
  if (s == 0x12345) sink[0] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int OP>
void run(const char *name, long long *cyc, uint32_t *sink) {
  const int iters = 20000, warps = 16;
  for (int rep = 0; rep < 2; ++rep) {
    k<OP><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
    cudaDeviceSynchronize();
  }
  printf("%-44s %.2f cycles per loop slot per SMSP\n", name, (double)cyc[0] / ((double)iters * 8 * (warps / 4)));
}
int main() {
  long long *cyc; uint32_t *sink;
  cudaMallocManaged(&cyc, 8); cudaMalloc(&sink, 4);
  run<0>("FFMA (denormal-range operands)", cyc, sink);
  run<1>("FFMA.SAT", cyc, sink);
  run<2>("FFMA2", cyc, sink);
  run<17>("FMUL2", cyc, sink);
  run<18>("FADD.RZ", cyc, sink);
  run<3>("IMAD", cyc, sink);
  run<4>("VIMNMX (max.s32)", cyc, sink);
  run<11>("VIMNMX.U16x2", cyc, sink);
  run<9>("FMNMX", cyc, sink);
  run<5>("SHF", cyc, sink);
  run<6>("IADD", cyc, sink);
  run<7>("IDP.4A", cyc, sink);
  run<8>("PRMT (register selector)", cyc, sink);
  run<12>("I2IP (cvt.pack.sat)", cyc, sink);
  run<13>("LOP3", cyc, sink);
  run<10>("LDS.32 (dependent address)", cyc, sink);
  run<14>("FFMA2 + SHF (two instructions per slot)", cyc, sink);
  run<15>("FFMA2 + IMAD (two instructions per slot)", cyc, sink);
  run<16>("SHF + IMAD (two instructions per slot)", cyc, sink);
  return 0;
}
