// Shared helpers for the sm_100a lifter kernels: error plumbing, PTX wrappers for mbarrier / TMA /
// tcgen05 (TMEM + UMMA), 16-bit conversions.  Everything here is inline PTX written for
// -gencode arch=compute_100a,code=sm_100a; there is no other architecture path.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/lifter_b200.h"

namespace lifter {

// ------------------------------------------------------------------------------------------------
// error plumbing (host): negative return codes + thread-local message, see include/lifter_b200.h
// ------------------------------------------------------------------------------------------------
// (the LIFTER_OK / LIFTER_ERR_* codes are the macros of the public header)

void set_error(const char* fmt, ...);
const char* get_error();

#define LIFTER_CHECK_CUDA(expr)                                                                   \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            ::lifter::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                                __LINE__);                                                        \
            return LIFTER_ERR_CUDA;                                                     \
        }                                                                                         \
    } while (0)

#define LIFTER_REQUIRE(cond, ...)                 \
    do {                                          \
        if (!(cond)) {                            \
            ::lifter::set_error(__VA_ARGS__);     \
            return LIFTER_ERR_INVALID;       \
        }                                         \
    } while (0)

#define LIFTER_PROPAGATE(expr)        \
    do {                              \
        int _r = (expr);              \
        if (_r != 0) return _r;       \
    } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int num_sms();   // cached SM count of the current device
// cudaFuncSetAttribute is PER DEVICE: a launcher keeps one PerDeviceOnce per kernel and configures the kernel the first
// time it runs on each device (a process-wide flag left the second GPU of a process unconfigured: "invalid argument").
struct PerDeviceOnce {
    bool done[64] = {};
    bool* slot() {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
        return &done[dev];
    }
};
bool pdl_enabled();   // programmatic dependent launch on every kernel (LIFTER_NO_PDL=1 turns it off)
long long launch_count();           // kernel launches issued by this library so far (captured launches count once)
long long dp_spin_limit_clocks();   // give-up time of the cross-GPU waits in SM clocks (0 = never); LIFTER_DP_TIMEOUT_S

// 16-bit operand formats (UMMA a_format/b_format values for kind::f16)
enum : int { FMT_F16 = 0, FMT_BF16 = 1 };

#ifdef __CUDACC__
// Launch with programmatic stream serialization: the kernel may be scheduled while its predecessor in the stream
// drains; every kernel of this library calls griddep_wait() before it touches global memory and
// griddep_launch() first thing, so launch latency and prologues overlap the previous kernel's tail.
// debug trace (lifter_trace_begin / lifter_trace_end): a CUDA event after every launch, read back as per-kernel times
void trace_mark(const char* name, cudaStream_t st);
int trace_begin(cudaStream_t st);
int trace_end(char* buf, int len);

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(const char* name, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    trace_mark(name, st);
    return e;
}

// ------------------------------------------------------------------------------------------------
// device helpers
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t lane_id() {
    uint32_t l;
    asm volatile("mov.u32 %0, %%laneid;" : "=r"(l));
    return l;
}

// one lane of a converged warp (nvcc knows the region is single-threaded: no waterfall loops around UR moves)
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---- mbarrier ----------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// arrive on the barrier at the same smem offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t cta) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar), "r"(cta));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
// Bounded wait: a lost arrival traps after ~4 s instead of hanging the GPU (a hung box is a strike).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    uint32_t spins = 0;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        if ((++spins & 0x3FF) == 0) {
            long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 8000000000LL) __trap();
        }
    }
}

// ---- TMA ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// 2D tile load global -> shared, completion on an mbarrier (bytes)
__device__ __forceinline__ void tma_load_2d(const void* tmap, uint32_t smem_dst, uint32_t bar, int c0,
                                            int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// same, delivered to the same smem offset (and mbarrier offset) of every CTA in cta_mask
__device__ __forceinline__ void tma_load_2d_multicast(const void* tmap, uint32_t smem_dst, uint32_t bar, int c0,
                                                      int c1, uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "h"(cta_mask)
        : "memory");
}
// 2-SM (cta_group::2) load: lands in THIS CTA's smem, completes bytes on the barrier at the same offset in the
// pair's leader CTA (bit 24 of a shared::cluster address selects the CTA of the pair)
__device__ __forceinline__ void tma_load_2d_2sm(const void* tmap, uint32_t smem_dst, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_dst), "l"(tmap), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1)
        : "memory");
}
// 2D tile store shared -> global (bulk async group)
__device__ __forceinline__ void tma_store_2d(const void* tmap, uint32_t smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tmap),
                 "r"(smem_src), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// generic-proxy smem writes -> visible to the async proxy (TMA store source)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- tcgen05 / TMEM ----------------------------------------------------------------------------
__device__ __forceinline__ void tcgen05_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_result_addr) {   // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result_addr),
                 "n"(COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {             // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t smem_result_addr) {   // same warp id in both CTAs of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result_addr),
                 "n"(COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
// 2-SM MMA: the pair's leader issues; A rows and B rows are split over the two CTAs' smem (same offsets)
__device__ __forceinline__ void umma_f16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar, uint16_t cta_mask) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(bar), "h"(cta_mask)
        : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; one thread issues for the CTA
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all previously issued UMMAs of this thread -> one arrival on the mbarrier when they complete
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                 : "memory");
}
// same, arriving on the barrier at this smem offset in every CTA of cta_mask
__device__ __forceinline__ void umma_commit_multicast(uint32_t bar, uint16_t cta_mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(bar), "h"(cta_mask)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {     // every thread of every CTA in the cluster
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets row (lane base + t)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float (&v)[32]) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
          "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
          "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// ---- one-shot grid barrier -----------------------------------------------------------------------
// Every block of a 1-D grid arrives once on `counter`, which an earlier kernel of the same stream left at zero (the
// first kernel of a forward / backward pass zeroes the pass's barrier words).  Needs every block of the grid resident at
// the same time: the launchers size such grids from the occupancy query, and nothing that is resident beside them waits
// on them.  Release / acquire at gpu scope orders the blocks' global writes before the readers' loads; data written by
// other blocks of the same launch must still be read past L1 (ld.global.cg / __ldcg).
constexpr int GRID_BARRIER_WORDS = 32;      // barrier words reserved per pass (layers x barriers per layer)
__device__ __forceinline__ void grid_barrier_once(unsigned* counter, unsigned nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned seen;
        asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(seen) : "l"(counter) : "memory");
        long long t0 = 0;
        unsigned spins = 0;
        while (true) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
            if (seen >= nblocks) break;
            if ((++spins & 0x3FF) == 0) {                    // a grid that cannot become co-resident traps, no hang
                const long long now = clock64();
                if (t0 == 0) t0 = now;
                else if (now - t0 > 8000000000LL) __trap();
            }
        }
    }
    __syncthreads();
}

// ---- 16-bit packing ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack16(float lo, float hi, int fmt) {
    if (fmt == FMT_BF16) {
        __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&t);
    }
    __half2 t = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack16(uint32_t w, int fmt) {
    if (fmt == FMT_BF16) return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
    return __half22float2(*reinterpret_cast<__half2*>(&w));
}
__device__ __forceinline__ uint16_t cvt16(float v, int fmt) {
    if (fmt == FMT_BF16) {
        __nv_bfloat16 t = __float2bfloat16_rn(v);
        return *reinterpret_cast<uint16_t*>(&t);
    }
    __half t = __float2half_rn(v);
    return *reinterpret_cast<uint16_t*>(&t);
}
__device__ __forceinline__ float cvt16f(uint16_t w, int fmt) {
    if (fmt == FMT_BF16) return __bfloat162float(*reinterpret_cast<__nv_bfloat16*>(&w));
    return __half2float(*reinterpret_cast<__half*>(&w));
}
#endif  // __CUDACC__

}  // namespace lifter
