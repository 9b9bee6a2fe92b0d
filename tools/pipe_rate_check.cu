// Developer check (GPU box): issue cost per SMSP of the epilogue's instructions -- cvt.pack.sat (I2IP), FFMA.RM on
// denormal-range operands, LOP3, IMAD.HI -- as cycles per warp-instruction with 4 warps per SMSP and 8 independent chains.
#include <cstdio>
#include <cstdint>
template <int OP>
__global__ void k(int iters, long long *cyc, uint32_t *sink, uint32_t seed, float fb, float fc) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 8 + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]), "r"(seed));
      if (OP == 1) a[i] = __float_as_uint(__fmaf_rd(__uint_as_float(a[i] & 0xFFFFFFu), fb, fc));
      if (OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]), "r"(seed));
      if (OP == 3) a[i] = __umulhi(a[i], seed) + 1;
      if (OP == 4) a[i] = __float_as_uint(__fmaf_rd(__uint_as_float(a[i] | 0x3f800000u), 1.0001f, 0.5f));   // normal-range FFMA
      if (OP == 5) asm volatile("shr.u32 %0, %1, %0;" : "+r"(a[i]) : "r"(seed));
      if (OP == 6) asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
      if (OP == 7 && (i & 1) == 0) {   // FFMA2.RM on denormal-range pairs: 4 instructions per 8 elements
        unsigned long long v, bb, cc;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(v) : "r"(a[i] & 0xFFFFFFu), "r"(a[i + 1] & 0xFFFFFFu));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(fb));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(fc));
        asm volatile("fma.rm.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(bb), "l"(cc));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=r"(a[i]), "=r"(a[i + 1]) : "l"(v));
      }
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= a[i];
  if (s == 0x12345) sink[0] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  long long *cyc; uint32_t *sink;
  cudaMallocManaged(&cyc, 8); cudaMalloc(&sink, 4);
  const char *names[] = {"cvt.pack.sat (I2IP)", "FFMA.RM denormal range (AND+FFMA)", "LOP3", "IMAD.HI(+IADD)", "FFMA.RM normal range (OR+FFMA)", "SHF", "PRMT", "FFMA2.RM denormal range (2 AND + FFMA2 per PAIR; per-element figure = this/1)"};
  const int iters = 20000, warps = 16;
  for (int op = 0; op < 8; ++op) {
    for (int rep = 0; rep < 2; ++rep) {
      if (op == 0) k<0><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 1) k<1><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 2) k<2><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 3) k<3><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 4) k<4><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 5) k<5><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 6) k<6><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      if (op == 7) k<7><<<148, warps * 32>>>(iters, cyc, sink, 7, 1e-3f, 1e-40f);
      cudaDeviceSynchronize();
    }
    double per = (double)cyc[0] / ((double)iters * 8 * (warps / 4));
    printf("%-36s %.2f cycles per warp-instruction per SMSP (loop body, incl. helper ops)\n", names[op], per);
  }
  return 0;
}
