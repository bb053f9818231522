// Diagnostic micro-benchmarks of the primitives the fused kernel is built from: how long a
// gpu-/system-scope fence and a flag store to a peer take after a burst of local or remote writes.
// Results (benchmarks/probe_fence.py) are recorded under profiles/; nothing here is on a hot path.
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "aar_b200.h"
#include "device_utils.cuh"

namespace aar {
int fail(int code, const std::string& what);

// One CTA = one probe.  stamps[cta][8]: start, burst issued, after fence.gpu, after fence.sys,
// after relaxed flag store, after second fence.sys (flag acked), after st.release.sys.
__global__ void __launch_bounds__(512, 1)
aar_probe_fence_kernel(uint4* burst_dst, int64_t burst_units_per_cta, uint32_t* flag_a, uint32_t* flag_b, uint32_t value,
                       uint64_t* stamps) {
  uint64_t* st = stamps + static_cast<int64_t>(blockIdx.x) * 8;
  if (threadIdx.x == 0) st[0] = globaltimer_ns();
  uint4* d = burst_dst + static_cast<int64_t>(blockIdx.x) * burst_units_per_cta;
  const uint4 v = make_uint4(value, value, value, value);
  for (int64_t i = threadIdx.x; i < burst_units_per_cta; i += blockDim.x) d[i] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    st[1] = globaltimer_ns();
    __threadfence();
    st[2] = globaltimer_ns();
    __threadfence_system();
    st[3] = globaltimer_ns();
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(flag_a + blockIdx.x), "r"(value) : "memory");
    st[4] = globaltimer_ns();
    __threadfence_system();
    st[5] = globaltimer_ns();
    st_release_sys(flag_b + blockIdx.x, value);
    st[6] = globaltimer_ns();
  }
}
}  // namespace aar

extern "C" int aar_probe_fence(void* burst_dst, int64_t burst_bytes_per_cta, void* flag_a, void* flag_b, uint32_t value,
                               int ctas, void* stamps, void* stream) {
  aar::aar_probe_fence_kernel<<<ctas, 512, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint4*>(burst_dst), burst_bytes_per_cta / 16, static_cast<uint32_t*>(flag_a),
      static_cast<uint32_t*>(flag_b), value, static_cast<uint64_t*>(stamps));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar::fail(AAR_ECUDA, std::string("probe launch: ") + cudaGetErrorString(e));
  return 1;
}
