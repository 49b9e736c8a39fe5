// Does lg2.approx(eps) * ln2 reproduce logf(eps) bit-for-bit?  (decides whether the epilogue needs its select)
#include <cstdio>
#include <cmath>
#include <cstring>
__global__ void k(float eps, float* out) {
  float l2;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(0.0f + eps));
  out[0] = l2 * 0.69314718055994530942f;
  out[1] = logf(eps);
  out[2] = __logf(eps);
}
int main() {
  float* d; cudaMalloc(&d, 12);
  float epss[4] = {1e-12f, 1e-10f, 1e-6f, 1e-5f};
  for (float e : epss) {
    k<<<1, 1>>>(e, d);
    float h[3]; cudaMemcpy(h, d, 12, cudaMemcpyDeviceToHost);
    float host = logf(e);
    unsigned u[4]; memcpy(u, h, 12); memcpy(u + 3, &host, 4);
    printf("eps %g: fast %.9g (%08x) logf_dev %.9g (%08x) __logf %.9g (%08x) logf_host %.9g (%08x)\n", e, h[0], u[0], h[1], u[1], h[2], u[2], host, u[3]);
  }
  return 0;
}
