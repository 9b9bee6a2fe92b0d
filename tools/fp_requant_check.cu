// Developer check (run on the GPU box): the denormal-range FFMA requantisation used by the fused kernel's RQM_FP mode is
// bit-identical to the integer form  sat_u8(((acc*mult + rnd) >> shift) + add + 128)  for every accumulator value.
//   bits = INIT + acc  (0 < bits < 2^24: as an f32 this is exactly bits * 2^-149)
//   r    = fma.rm(bits_as_float, mult * 2^-shift, C * 2^-149),  INIT*mult == rnd (mod 2^shift),
//   C    = (rnd - INIT*mult) / 2^shift + add + 128
//   out  = cvt.pack.sat.u8.s32(r_bits)
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <cstdlib>
__global__ void kcheck(int init, float b, float c, int mult, int shift, long long rnd, int add, int lo, int hi, unsigned long long *bad) {
  for (long long acc = lo + (long long)(blockIdx.x * blockDim.x + threadIdx.x); acc <= hi; acc += (long long)gridDim.x * blockDim.x) {
    int bits = init + (int)acc;
    float r = __fmaf_rd(__int_as_float(bits), b, c);
    int rb = __float_as_int(r);
    uint32_t packed;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(packed) : "r"(0), "r"(rb), "r"(0));
    int got = packed & 0xFF;
    long long y = ((acc * mult + rnd) >> shift) + add + 128;
    int want = y < 0 ? 0 : (y > 255 ? 255 : (int)y);
    if (got != want) atomicAdd(bad, 1ULL);
  }
}
// ITAmax distance + 1: d = (m + 128) - z,  z = ((acc*mult + rnd) >> shift) + add + 128, from one FFMA.RP; mp = m + 128
__global__ void kdist(int init, float b, int negc, int mp, int mult, int shift, long long rnd, int add, int lo, int hi, unsigned long long *bad) {
  const float c2 = __int_as_float(mp + 1 + negc), nfb = -b;   // d + 1 (never in (-1, 0])
  for (long long acc = lo + (long long)(blockIdx.x * blockDim.x + threadIdx.x); acc <= hi; acc += (long long)gridDim.x * blockDim.x) {
    int bits = init + (int)acc;
    unsigned got = (unsigned)__float_as_int(__fmaf_ru(__int_as_float(bits), nfb, c2));
    long long z = ((acc * mult + rnd) >> shift) + add + 128;
    if (z > mp) continue;                      // not a fast row
    unsigned long long want = (unsigned long long)(mp - z) + 1;
    if (want < 64 ? got != want : got < 64) atomicAdd(bad, 1ULL);
  }
}
// RQM_TORCH_FP2 (two packed FFMA2 per element pair, power-of-two offset) against the plain fp32 form
//   want = sat_u8(rne(fp32(acc) * eff) + zp + 128)
__global__ void ktorch2(int init, float eff, int zp, int lo, int hi, unsigned long long *bad) {
  const float b1 = ldexpf(eff, 85), c1 = -ldexpf(eff * (float)init, -64), fc = __int_as_float(zp + 128);
  unsigned long long B1, C1, M, FC;
  asm("mov.b64 %0, {%1, %1};" : "=l"(B1) : "f"(b1));
  asm("mov.b64 %0, {%1, %1};" : "=l"(C1) : "f"(c1));
  asm("mov.b64 %0, {%1, %1};" : "=l"(M) : "r"(0x15000000));
  asm("mov.b64 %0, {%1, %1};" : "=l"(FC) : "f"(fc));
  for (long long acc = lo + 2 * (long long)(blockIdx.x * blockDim.x + threadIdx.x); acc + 1 <= hi; acc += 2LL * gridDim.x * blockDim.x) {
    unsigned long long a, t, d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "r"(init + (int)acc), "r"(init + (int)acc + 1));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t) : "l"(a), "l"(B1), "l"(C1));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(t), "l"(M), "l"(FC));
    int r0, r1;
    asm("mov.b64 {%0, %1}, %2;" : "=r"(r0), "=r"(r1) : "l"(d));
    for (int k = 0; k < 2; ++k) {
      uint32_t packed;
      asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(packed) : "r"(0), "r"(k ? r1 : r0), "r"(0));
      int got = packed & 0xFF;
      long long y = (long long)__float2int_rn(__fmul_rn(__int2float_rn((int)acc + k), eff)) + zp + 128;
      int want = y < 0 ? 0 : (y > 255 ? 255 : (int)y);
      if (got != want) atomicAdd(bad, 1ULL);
    }
  }
}
// RQM_FP with FFMA2.RM: same check as kcheck, two elements per instruction
__global__ void kcheck2(int init, float b, float c, int mult, int shift, long long rnd, int add, int lo, int hi, unsigned long long *bad) {
  unsigned long long Bp, Cp;
  asm("mov.b64 %0, {%1, %1};" : "=l"(Bp) : "f"(b));
  asm("mov.b64 %0, {%1, %1};" : "=l"(Cp) : "f"(c));
  for (long long acc = lo + 2 * (long long)(blockIdx.x * blockDim.x + threadIdx.x); acc + 1 <= hi; acc += 2LL * gridDim.x * blockDim.x) {
    unsigned long long a, d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "r"(init + (int)acc), "r"(init + (int)acc + 1));
    asm("fma.rm.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(Bp), "l"(Cp));
    int r0, r1;
    asm("mov.b64 {%0, %1}, %2;" : "=r"(r0), "=r"(r1) : "l"(d));
    for (int k = 0; k < 2; ++k) {
      uint32_t packed;
      asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(packed) : "r"(0), "r"(k ? r1 : r0), "r"(0));
      int got = packed & 0xFF;
      long long y = (((acc + k) * mult + rnd) >> shift) + add + 128;
      int want = y < 0 ? 0 : (y > 255 ? 255 : (int)y);
      if (got != want) atomicAdd(bad, 1ULL);
    }
  }
}
static long long inv_odd(long long a, int bits) {  // inverse of odd a modulo 2^bits
  long long x = 1;
  for (int i = 0; i < 6; ++i) x = x * (2 - a * x);
  return bits >= 63 ? x : (x & ((1LL << bits) - 1));
}
int main() {
  unsigned long long *bad;
  cudaMallocManaged(&bad, 8);
  int fails = 0;
  srand(1);
  for (int trial = 0; trial < 60; ++trial) {
    int mult = 1 + rand() % 255, shift = 6 + rand() % 17, add = (rand() % 17) - 8;
    if (trial == 0) { mult = 128; shift = 9; add = 0; }
    if (trial == 1) { mult = 255; shift = 22; add = 0; }
    long long rnd = (trial % 5 == 4) ? 0 : (1LL << (shift - 1));
    int t = 0;
    while (((mult >> t) & 1) == 0) ++t;
    if (t > shift - 1 && rnd) continue;
    int mb = shift - t;                                  // modulus 2^mb
    long long mo = mult >> t;
    long long target = ((rnd >> t) % (1LL << mb));
    long long k = (target * inv_odd(mo, mb)) & ((1LL << mb) - 1);    // INIT == k (mod 2^mb)
    long long init = (1LL << 23) - (((1LL << 23) - k) & ((1LL << mb) - 1));  // largest value <= 2^23 congruent to k
    if ((init * mult - rnd) % (1LL << shift)) { printf("congruence broken\n"); return 1; }
    long long C = (rnd - init * mult) / (1LL << shift) + add + 128;
    float b = ldexpf((float)mult, -shift), c = ldexpf((float)C, -149);
    if ((long long)ldexpf(c, 149) != C) { printf("C not representable\n"); continue; }
    int lim = (1 << 22) + 100000;
    *bad = 0;
    kcheck<<<296, 256>>>((int)init, b, c, mult, shift, rnd, add, -lim, lim, bad);
    cudaDeviceSynchronize();
    printf("mult %3d shift %2d add %2d rnd %lld init %lld C %lld: %llu mismatches\n", mult, shift, add, rnd, init, C, *bad);
    fails += *bad != 0;
    *bad = 0;
    kcheck2<<<296, 256>>>((int)init, b, c, mult, shift, rnd, add, -lim, lim, bad);
    cudaDeviceSynchronize();
    printf("   FFMA2.RM: %llu mismatches\n", *bad);
    fails += *bad != 0;
    if (C < 0) {
      *bad = 0;
      kdist<<<296, 256>>>((int)init, b, (int)-C, 200, mult, shift, rnd, add, -lim, lim, bad);
      cudaDeviceSynchronize();
      printf("   dist: %llu mismatches\n", *bad);
      fails += *bad != 0;
    }
  }
  // fp32 requant form in two FFMA2 (power-of-two offsets 2^20 .. 2^22), random and awkward scales, zero-points 0 and 6
  for (int trial = 0; trial < 24; ++trial) {
    float eff = ldexpf(0.5f + (rand() % 8388608) / 16777216.0f, -(2 + rand() % 14));
    if (trial == 0) eff = 0.0013240111f;
    if (trial == 1) eff = ldexpf(1.f, -7);
    if (trial == 2) eff = 0.99999994f * ldexpf(1.f, -5);
    const int zp = (trial % 3 == 2) ? 6 : 0;
    const int init = 1 << (20 + trial % 3);
    const int lim = init - 2;
    *bad = 0;
    ktorch2<<<296, 256>>>(init, eff, zp, -lim, lim, bad);
    cudaDeviceSynchronize();
    printf("torch form, 2 x FFMA2: eff %.9g zp %d init 2^%d: %llu mismatches\n", eff, zp, 20 + trial % 3, *bad);
    fails += *bad != 0;
  }
  printf(fails ? "FAILED\n" : "ALL EXACT\n");
  return fails != 0;
}
