// Exhaustive on-device checks of the arithmetic shortcuts the fused kernel
// takes (dpk_kernels.cuh: div_by_scalar, ElemTraits<bf16>::sqrt_for_store, and the
// bf16 conversions: rnd() = one packed conversion with a zero partner, cvt_lo/hi =
// mixed-precision add of -0.0).
#include "dpk_internal.h"

namespace dpk {

// The dividend in the kernel is always a square root that was just stored:
// x = sqrt.rn(v) for fp32 state, x = bf16(sqrt(v)) for bf16 state.  For every
// fp32 bit pattern v (all 2^32, negatives and NaN included) both forms of x go
// through div_by_scalar(x, c, RN(1/c)) and must equal div.rn(x, c).
// (Outside that domain -- subnormal x, where the remainder underflows -- the
// shortcut is NOT exact; the kernel never produces such x: a square root of an
// fp32 number is 0 or at least 2^-75.)
__global__ void selftest_div_kernel(float c, float rc, unsigned long long* out) {
  unsigned long long bad = 0, bad_cvt = 0;
  const unsigned long long n = 1ull << 32;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const float v = __uint_as_float((unsigned)i);
    float xs[4] = {__fsqrt_rn(v), sqrt_approx(__uint_as_float((unsigned)i & 0xffff0000u)), 0.f, 0.f};
    xs[1] = ElemTraits<__nv_bfloat16>::rnd(xs[1]);
    {  // rnd(v) must be the IEEE round-to-nearest-even conversion, as an fp32 value
      const float want = __bfloat162float(__float2bfloat16_rn(v));
      const float got = ElemTraits<__nv_bfloat16>::rnd(v);
      if (!((want != want) && (got != got)) && __float_as_uint(want) != __float_as_uint(got)) ++bad_cvt;
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float want = __fdiv_rn(xs[k], c);
      const float got = div_by_scalar(xs[k], c, rc);
      const bool both_nan = (want != want) && (got != got);
      // sqrt(-0.0) = -0.0 comes back as +0.0: exp_avg_sq is never -0.0, and the
      // quotient's only use is `+ eps`, which erases the sign of a zero
      const bool both_zero = (want == 0.f) && (got == 0.f);
      if (!both_nan && !both_zero && __float_as_uint(want) != __float_as_uint(got)) ++bad;
    }
  }
  if (bad) atomicAdd(out, bad);
  if (bad_cvt) atomicAdd(out + 1, bad_cvt);
}

// every bf16 bit pattern v: bf16(sqrt.approx(v)) must equal bf16(sqrt.rn(v))
__global__ void selftest_sqrt_kernel(unsigned long long* out) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 65536u) return;
  const float v = __uint_as_float(i << 16);
  float a[4] = {__fsqrt_rn(v), ElemTraits<__nv_bfloat16>::sqrt_for_store(v), 0.f, 0.f};
  ElemTraits<__nv_bfloat16>::rnd4(a);
  const bool both_nan = (a[0] != a[0]) && (a[1] != a[1]);
  if (!both_nan && __float_as_uint(a[0]) != __float_as_uint(a[1])) atomicAdd(out + 1, 1ull);
  // unpacking either half of a packed pair is exact (NaN stays NaN), whatever the other half holds
  const unsigned other = (i * 2654435761u) >> 16;
  const float lo = ElemTraits<__nv_bfloat16>::cvt_lo((other << 16) | i), hi = ElemTraits<__nv_bfloat16>::cvt_hi((i << 16) | other);
  const bool is_nan = (v != v);
  if (is_nan ? !((lo != lo) && (hi != hi)) : (__float_as_uint(lo) != (i << 16) || __float_as_uint(hi) != (i << 16)))
    atomicAdd(out + 1, 1ull);
}

}  // namespace dpk

using namespace dpk;

extern "C" int dpk_selftest_arith(double divisor, unsigned long long* d_out2, void* stream) {
  if (!d_out2 || !(divisor > 0.0)) { set_error("dpk_selftest_arith: bad argument"); return DPK_E_BADARG; }
  cudaStream_t st = (cudaStream_t)stream;
  const float c = (float)divisor;
  const float rc = (float)(1.0 / (double)c);
  DPK_CUDA(cudaMemsetAsync(d_out2, 0, 2 * sizeof(unsigned long long), st));
  selftest_div_kernel<<<148 * 8, 256, 0, st>>>(c, rc, d_out2);
  DPK_CUDA(cudaGetLastError());
  selftest_sqrt_kernel<<<256, 256, 0, st>>>(d_out2);
  DPK_CUDA(cudaGetLastError());
  return 0;
}
