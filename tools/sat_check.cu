// Developer check (GPU box): what fma.{rn,rp}.sat.f32 returns for results that are -0.0 or negative denormals
// (the ITAmax distance trick wants +0.0, bit pattern 0, in both cases).
#include <cstdio>
#include <cstdint>
__global__ void k(uint32_t *out) {
  const float d3 = __uint_as_float(3u), d1 = __uint_as_float(1u), d5 = __uint_as_float(5u), d7 = __uint_as_float(7u);
  float r;
  asm volatile("fma.rp.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d3), "f"(-0.1f), "f"(0.0f)); out[0] = __float_as_uint(r);      // -0.3 ulp -> -0.0
  asm volatile("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d1), "f"(-0.4f), "f"(0.0f)); out[1] = __float_as_uint(r);      // -0.4 ulp -> -0.0
  asm volatile("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d5), "f"(-1.0f), "f"(0.0f)); out[2] = __float_as_uint(r);      // -5 ulp
  asm volatile("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d5), "f"(-1.0f), "f"(d7)); out[3] = __float_as_uint(r);        // +2 ulp
  asm volatile("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d5), "f"(-1.0f), "f"(d5)); out[4] = __float_as_uint(r);        // exact 0
  asm volatile("fma.rp.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d3), "f"(-0.1f), "f"(0.0f)); out[5] = __float_as_uint(r);          // no sat: -0.0 expected
  asm volatile("fma.rm.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(d3), "f"(0.5f), "f"(d1)); out[6] = __float_as_uint(r);         // floor(2.5) = 2
  asm volatile("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(1e30f), "f"(1e-30f * 1e-35f), "f"(-d7)); out[7] = __float_as_uint(r);   // scaled normal * tiny - 7 ulp
}
int main() {
  uint32_t *o; cudaMallocManaged(&o, 64);
  k<<<1, 1>>>(o); cudaDeviceSynchronize();
  const char *n[] = {"rp.sat(-0.3ulp)", "rn.sat(-0.4ulp)", "rn.sat(-5ulp)", "rn.sat(+2ulp)", "rn.sat(exact 0)", "rp(-0.3ulp) no sat", "rm.sat(2.5ulp)", "rn.sat(normal*tiny - 7ulp)"};
  for (int i = 0; i < 8; ++i) printf("%-28s -> 0x%08x\n", n[i], o[i]);
  return 0;
}
