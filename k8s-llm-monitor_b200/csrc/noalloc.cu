// klm_dev_node_best_static: the node_best_reduce step for callers that may not
// allocate device memory.
//
// Who needs this: hosts that are not allowed to allocate GPU memory themselves.
// Examples are a cgo shim loaded into a process whose HBM was handed out in full
// to a pool at start-up (an MPS client running under a fixed memory reservation, or
// a sidecar sharing the context of the component that owns the device), and a
// scheduler reconcile loop that bans cudaMalloc on its hot path because the call
// synchronises the whole device and stalls every other stream. klm_dev_node_best
// makes such a caller find 33 KB of scratch somewhere; this entry point brings its
// own.
//
// The scratch (a completion ticket and one partial winner per block) lives in a
// file-scope __device__ array of fixed size. Every GPU that loads libklm.so holds
// those 19 200 bytes until the process exits. The fixed size bounds the problem:
// at most 4 849 664 nodes per call (klm_dev_node_best_static_max_nodes()); larger tables are refused with
// KLM_ELIMIT (use klm_dev_node_best with caller scratch for those).
//
// Sharing: all calls on one device use the same array. It keeps whatever the last
// call left, so the ticket is cleared on the stream before every launch, and two
// calls whose kernels could overlap on the GPU (different streams without an
// ordering between them, or two host threads) would race on it: serialise them.
//
// Set-up: the first call on a device looks the array's device address up
// (cudaGetSymbolAddress, with that device current) and remembers it. Make that
// first call before starting any stream capture.
// The remembered address belongs to the device's primary context: it does not
// survive cudaDeviceReset() (or anything else that destroys that context) -- a
// process that resets a device must not call this entry point on it afterwards.
//
// Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
#include <mutex>

#include "common.cuh"

namespace klm {

constexpr int kSbThreads = 256;
constexpr int kSbNodesPerBlock = 4096;                 // one trip of 16 nodes per thread
constexpr int kSbMaxBlocks = 1184;                     // 148 SMs x 8
constexpr int64_t kSbMaxNodes = (int64_t)kSbMaxBlocks * kSbNodesPerBlock;  // 4 849 664
constexpr int kSbScratchBytes = 256 + kSbMaxBlocks * 16;                   // 19 200

// [0, 256): ticket (u32). Then f64 partial batteries [kSbMaxBlocks], then i64 partial
// indices [kSbMaxBlocks]. In HOST code this name is only a handle for
// cudaGetSymbolAddress, never an address to hand to a kernel.
static __device__ __align__(256) unsigned char g_static_scratch[kSbScratchBytes];

struct SbBest {
  double bat;
  int64_t idx;  // -1: none
};

__device__ __forceinline__ SbBest sb_better(SbBest a, SbBest b) {
  if (b.idx < 0) return a;
  if (a.idx < 0) return b;
  if (b.bat > a.bat || (b.bat == a.bat && b.idx < a.idx)) return b;
  return a;
}

__device__ __forceinline__ SbBest sb_block_best(SbBest v) {
  __shared__ double s_bat[kSbThreads / 32];
  __shared__ int64_t s_idx[kSbThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    SbBest w;
    w.bat = __shfl_xor_sync(0xffffffffu, v.bat, o);
    w.idx = __shfl_xor_sync(0xffffffffu, v.idx, o);
    v = sb_better(v, w);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // the arrays may still be read from an earlier call in this block
  if (lane == 0) {
    s_bat[warp] = v.bat;
    s_idx[warp] = v.idx;
  }
  __syncthreads();
  if (warp == 0) {
    SbBest w;
    w.bat = lane < kSbThreads / 32 ? s_bat[lane] : 0.0;
    w.idx = lane < kSbThreads / 32 ? s_idx[lane] : -1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      SbBest u;
      u.bat = __shfl_xor_sync(0xffffffffu, w.bat, o);
      u.idx = __shfl_xor_sync(0xffffffffu, w.idx, o);
      w = sb_better(w, u);
    }
    v = w;
  }
  return v;  // valid in warp 0
}

// Same selection rule as node_best_kernel (schedule.cu): largest battery among nodes with
// (flags & 3) == 3, lowest index on ties — internal/scheduler/controller.go:190, :198 and
// :160-161 of the reference. One block per 4096 nodes, no grid-stride, so the number of
// partials is known from n alone. `scratch` is the device address of g_static_scratch.
__global__ void __launch_bounds__(kSbThreads)
node_best_static_kernel(const double* __restrict__ battery, const uint8_t* __restrict__ flags,
                        int64_t n, int64_t node_base, int64_t* __restrict__ best,
                        unsigned char* __restrict__ scratch) {
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* part_bat = reinterpret_cast<double*>(scratch + 256);
  int64_t* part_idx = reinterpret_cast<int64_t*>(scratch + 256 + kSbMaxBlocks * 8);

  SbBest me{0.0, -1};
  const int64_t lo = (int64_t)blockIdx.x * kSbNodesPerBlock;
  const int64_t hi = min(n, lo + kSbNodesPerBlock);
  for (int64_t i = lo + threadIdx.x; i < hi; i += kSbThreads) {
    const double v = battery[i];
    // a thread's indices grow, so a strict > keeps its first maximum
    if ((flags[i] & 3u) == 3u && (me.idx < 0 || v > me.bat)) {
      me.bat = v;
      me.idx = i;
    }
  }
  me = sb_block_best(me);

  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    part_bat[blockIdx.x] = me.bat;
    part_idx[blockIdx.x] = me.idx;
    __threadfence();
    s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  SbBest acc{0.0, -1};
  for (int b = threadIdx.x; b < (int)gridDim.x; b += kSbThreads) {
    SbBest p;
    p.bat = __ldcg(part_bat + b);
    p.idx = __ldcg(part_idx + b);
    acc = sb_better(acc, p);
  }
  acc = sb_block_best(acc);
  if (threadIdx.x == 0) {
    best[0] = acc.idx < 0 ? -1 : acc.idx + node_base;
    best[1] = __double_as_longlong(acc.idx < 0 ? 0.0 : acc.bat);
    *ticket = 0;
  }
}

// Device address of g_static_scratch, looked up once per device.
static int static_scratch_ptr(unsigned char** out) {
  static std::mutex mu;
  static unsigned char* ptrs[64] = {};
  int dev = 0;
  KLM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("klm_dev_node_best_static: device ordinal %d not supported", dev);
    return KLM_EINVAL;
  }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    unsigned char* p = nullptr;
    // the current device is `dev`, so this is that device's copy of the array
    KLM_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch));
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return KLM_OK;
}

}  // namespace klm

using namespace klm;

extern "C" int64_t klm_dev_node_best_static_max_nodes(void) { return kSbMaxNodes; }

extern "C" int klm_dev_node_best_static(const double* battery, const uint8_t* flags,
                                        int64_t n_nodes, int64_t node_base, int64_t* best,
                                        void* stream) {
  if (n_nodes < 0 || !best || (n_nodes > 0 && (!battery || !flags))) {
    set_error("klm_dev_node_best_static: bad argument");
    return KLM_EINVAL;
  }
  if (n_nodes > kSbMaxNodes) {
    set_error("klm_dev_node_best_static: %lld nodes exceed the built-in scratch (max %lld); "
              "use klm_dev_node_best with caller-provided scratch",
              (long long)n_nodes, (long long)kSbMaxNodes);
    return KLM_ELIMIT;
  }
  unsigned char* scratch = nullptr;
  const int rc = static_scratch_ptr(&scratch);
  if (rc != KLM_OK) return rc;
  cudaStream_t st = as_stream(stream);
  // the array keeps the previous call's bytes: the ticket must start at zero
  KLM_CUDA(cudaMemsetAsync(scratch, 0, 256, st));
  int blocks = (int)ceil_div(n_nodes, kSbNodesPerBlock);
  if (blocks < 1) blocks = 1;
  node_best_static_kernel<<<blocks, kSbThreads, 0, st>>>(battery, flags, n_nodes, node_base, best,
                                                         scratch);
  KLM_LAUNCH_CHECK("node_best_static_kernel");
  return KLM_OK;
}
