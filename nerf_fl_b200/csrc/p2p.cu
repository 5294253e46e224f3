// (e) Multi-GPU gradient exchange over NVLink / NVSwitch peer memory: the reference's DDP all-reduce (Lightning,
// train.py:235-236) as ONE kernel per rank that reads the other ranks' flat gradient buffers directly (CUDA IPC mappings,
// P2P loads) -- no NCCL call, no host round trip, capturable inside the training iteration's CUDA graph.
//
// One process per GPU.  Every rank owns one exchange region (nwb_p2p_alloc: plain cudaMalloc, exported with
// cudaIpcGetMemHandle and opened by every peer):
//     [P2PSignal 1 KB][flat gradient buffer g: n floats][reduced slice: ceil(n / world) floats (+ padding)]
// Two-shot all-reduce, every sum taken in rank order 0..W-1 so that all ranks end with BIT-IDENTICAL gradients:
//   barrier A   every rank's backward has finished (kernel boundary on its own stream) and it says so to all peers
//   phase 1     rank r reads slice r of every rank's g (W-1 of them over NVLink), sums, writes its own `reduced` buffer
//   barrier B   all slices are reduced -- and nobody reads my g any more
//   phase 2     every rank copies all W reduced slices (W-1 over NVLink) into its own g
// Bytes over NVLink per rank: 2 (W-1)/W n x 4 B (9.6 MB at W = 8 for the 1.38 M-float model) -- latency-bound.
// The cross-GPU barriers are flags in peer memory: rank q writes the launch's epoch into signal[r].arrive[phase][q] of
// every peer r and waits for its own row to fill.  Epochs only grow, so nothing is ever reset; a wait that does not
// complete within ~2 s traps (a launch error on this rank) instead of hanging the GPU.
// All ranks' kernels run at the same time on DIFFERENT GPUs (never two waiting kernels on one device).
#include "common.cuh"
#include <string.h>

namespace nwb {

constexpr int kP2PMaxWorld = 16;
constexpr int kP2PBlocks = 128;            // co-resident by construction (<= 148 SMs, one 1024-thread block each, no shared memory)
constexpr int kP2PThreads = 1024;          // 128 K threads x 16 B x W loads in flight: NVLink round trips are ~2-3 us

struct P2PSignal {
    unsigned int arrive[2][kP2PMaxWorld];  // written by the peers (system scope)
    unsigned int epoch;                    // launches completed by this rank (own writes only)
    unsigned int go[2];                    // block 0 -> the other blocks of this rank
    unsigned int pad0;
    unsigned long long stamp[6];           // %globaltimer of block 0 / thread 0 in the last launch: start, barrier A passed, phase 1
                                           // done on this rank, barrier B passed, phase 2 done (tools/p2p_check.py prints them)
    unsigned int pad[256 - 2 * kP2PMaxWorld - 4 - 12];
};
static_assert(sizeof(P2PSignal) == 1024, "signal block is 1 KB");

struct P2PArgs {
    int rank, world;
    long long n;                           // floats in g
    long long slice;                       // floats per slice (multiple of 4)
    P2PSignal* sig[kP2PMaxWorld];          // signal block of every rank (own entry = local pointer)
    float* g[kP2PMaxWorld];                // flat gradient buffer of every rank
    float* red[kP2PMaxWorld];              // reduced-slice buffer of every rank
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// peer data: never through a cache line an earlier launch may have left behind
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

template <typename F>
__device__ __forceinline__ void spin_until(F ready, const char* what) {
    const long long t0 = clock64();
    while (!ready()) {
        if (clock64() - t0 > 4000000000LL) {
            printf("nwb: p2p all-reduce timed out waiting for %s (block %d thread %d)\n", what, blockIdx.x, threadIdx.x);
            __trap();
        }
        __nanosleep(64);
    }
}

// cross-GPU barrier `phase` of the launch with epoch `e`, then release of this rank's other blocks
__device__ __forceinline__ void p2p_barrier(const P2PArgs& a, unsigned int e, int phase) {
    P2PSignal* mine = a.sig[a.rank];
    if (blockIdx.x == 0) {
        __syncthreads();                               // (phase 1 -> barrier B: this block's stores are issued)
        if (threadIdx.x < a.world) {
            const int q = threadIdx.x;
            __threadfence_system();
            st_release_sys(&a.sig[q]->arrive[phase][a.rank], e);
            spin_until([&] { return static_cast<int>(ld_acquire_sys(&mine->arrive[phase][q]) - e) >= 0; }, "a peer");
        }
        __syncthreads();
        if (threadIdx.x == 0) st_release_gpu(&mine->go[phase], e);
    } else {
        if (threadIdx.x == 0) spin_until([&] { return static_cast<int>(ld_acquire_gpu(&mine->go[phase]) - e) >= 0; }, "block 0");
        __syncthreads();
    }
}

// every block must have finished its phase-1 stores before block 0 tells the peers: count the blocks in
__global__ void __launch_bounds__(kP2PThreads, 1)
p2p_allreduce_kernel(const P2PArgs a, unsigned int* block_count, int gather) {
    P2PSignal* mine = a.sig[a.rank];
    const unsigned int e = mine->epoch + 1;            // same on every rank: all ranks launch the same sequence
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;

    const bool stamper = blockIdx.x == 0 && threadIdx.x == 0;
    if (stamper) mine->stamp[0] = globaltimer_ns();
    p2p_barrier(a, e, 0);
    if (stamper) mine->stamp[1] = globaltimer_ns();

    // ---- phase 1: my slice of everybody's gradients, summed in rank order ----
    const long long lo = a.slice * a.rank;
    const long long hi = lo + a.slice < a.n ? lo + a.slice : a.n;          // n and slice are multiples of 4
    for (long long i = lo + 4 * tid; i < hi; i += 4 * nth) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q0 = 0; q0 < a.world; q0 += 4) {       // four ranks' loads in flight at a time, added in rank order
            float4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (q0 + k < a.world) v[k] = ld_peer_f4(a.g[q0 + k] + i);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (q0 + k < a.world) {
                    if (q0 + k == 0) s = v[0];
                    else { s.x += v[k].x; s.y += v[k].y; s.z += v[k].z; s.w += v[k].w; }
                }
        }
        *reinterpret_cast<float4*>(a.red[a.rank] + (i - lo)) = s;
    }
    // all blocks of this rank are done with phase 1 before block 0 signals barrier B
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(block_count, 1u);
        if (blockIdx.x == 0) spin_until([&] { return ld_acquire_gpu(block_count) >= gridDim.x; }, "this rank's blocks");
    }
    if (stamper) mine->stamp[2] = globaltimer_ns();
    p2p_barrier(a, e, 1);
    if (stamper) mine->stamp[3] = globaltimer_ns();

    // ---- phase 2: gather every rank's reduced slice into my g (gather == 0: the consumer -- the optimizer kernel -- reads the
    // reduced slices from the peers itself, nwb_adam_pack_step_p2p) ----
    for (long long i = 4 * tid; gather && i < a.n; i += 4 * nth) {
        const int q = static_cast<int>(i / a.slice);
        const float4 v = ld_peer_f4(a.red[q] + (i - a.slice * q));
        *reinterpret_cast<float4*>(a.g[a.rank] + i) = v;
    }
    if (stamper) mine->stamp[4] = globaltimer_ns();
    // last block out resets the block counter and publishes the epoch (own memory, next launch on this stream reads it)
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int done = atomicAdd(block_count, 1u) + 1u;
        if (done == 2u * gridDim.x) {
            *block_count = 0u;
            mine->epoch = e;
        }
    }
}

}  // namespace nwb

using namespace nwb;

extern "C" size_t nwb_p2p_region_bytes(int64_t n_floats, int world) {
    if (n_floats <= 0 || world < 1 || world > kP2PMaxWorld) return 0;
    const int64_t n4 = (n_floats + 3) / 4 * 4;
    const int64_t slice = ((n4 + world - 1) / world + 3) / 4 * 4;
    return sizeof(P2PSignal) + static_cast<size_t>(n4) * 4 + static_cast<size_t>(slice) * 4 + 256;
}

extern "C" int nwb_p2p_alloc(size_t bytes, void** ptr) {
    NWB_REQUIRE(ptr != nullptr && bytes > 0, "nwb_p2p_alloc: bad argument");
    NWB_CUDA(cudaMalloc(ptr, bytes));                  // a plain allocation: exportable with cudaIpcGetMemHandle at offset 0
    NWB_CUDA(cudaMemset(*ptr, 0, bytes));
    return NWB_OK;
}
extern "C" int nwb_p2p_free(void* ptr) {
    if (ptr != nullptr) NWB_CUDA(cudaFree(ptr));
    return NWB_OK;
}
extern "C" int nwb_p2p_export(const void* ptr, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    NWB_REQUIRE(ptr != nullptr && handle != nullptr, "nwb_p2p_export: null pointer");
    cudaIpcMemHandle_t h;
    NWB_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
    memcpy(handle, &h, 64);
    return NWB_OK;
}
extern "C" int nwb_p2p_open(const unsigned char handle[64], void** ptr) {
    NWB_REQUIRE(ptr != nullptr && handle != nullptr, "nwb_p2p_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    NWB_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return NWB_OK;
}
extern "C" int nwb_p2p_close(void* ptr) {
    if (ptr != nullptr) NWB_CUDA(cudaIpcCloseMemHandle(ptr));
    return NWB_OK;
}

// regions[q] = base of rank q's exchange region as mapped in THIS process (own region: the local pointer);
// counter: 4 zeroed bytes of local device memory (per communicator)
static int p2p_launch(void* const* regions, int rank, int world, int64_t n_floats, void* counter, int gather, void* stream);

extern "C" int nwb_p2p_allreduce(void* const* regions, int rank, int world, int64_t n_floats, void* counter, void* stream) {
    return p2p_launch(regions, rank, world, n_floats, counter, 1, stream);
}
// first half only: on return (stream order) slice q of the summed gradient sits in rank q's reduced buffer
// (nwb_p2p_reduced_slice gives the addresses); the gradient buffers themselves are left as they were
extern "C" int nwb_p2p_reduce_scatter(void* const* regions, int rank, int world, int64_t n_floats, void* counter, void* stream) {
    return p2p_launch(regions, rank, world, n_floats, counter, 0, stream);
}
// layout query shared with the optimizer kernel: floats per slice, and where rank q's reduced slice starts inside its region
extern "C" int64_t nwb_p2p_slice_floats(int64_t n_floats, int world) {
    if (n_floats <= 0 || world < 1) return 0;
    const int64_t n4 = (n_floats + 3) / 4 * 4;
    return ((n4 + world - 1) / world + 3) / 4 * 4;
}
extern "C" size_t nwb_p2p_reduced_offset(int64_t n_floats) {
    return sizeof(P2PSignal) + static_cast<size_t>((n_floats + 3) / 4 * 4) * 4;
}

static int p2p_launch(void* const* regions, int rank, int world, int64_t n_floats, void* counter, int gather, void* stream) {
    NWB_REQUIRE(regions != nullptr && counter != nullptr, "nwb_p2p_allreduce: null pointer");
    NWB_REQUIRE(world >= 2 && world <= kP2PMaxWorld && rank >= 0 && rank < world, "nwb_p2p_allreduce: bad rank / world");
    NWB_REQUIRE(n_floats > 0, "nwb_p2p_allreduce: empty buffer");
    P2PArgs a{};
    a.rank = rank; a.world = world;
    a.n = (n_floats + 3) / 4 * 4;
    a.slice = ((a.n + world - 1) / world + 3) / 4 * 4;
    for (int q = 0; q < world; ++q) {
        NWB_REQUIRE(regions[q] != nullptr, "nwb_p2p_allreduce: region %d is null", q);
        uint8_t* base = static_cast<uint8_t*>(regions[q]);
        a.sig[q] = reinterpret_cast<P2PSignal*>(base);
        a.g[q] = reinterpret_cast<float*>(base + sizeof(P2PSignal));
        a.red[q] = a.g[q] + a.n;
    }
    int dev = 0;
    int rc = current_device_checked(&dev);
    if (rc != NWB_OK) return rc;
    p2p_allreduce_kernel<<<kP2PBlocks, kP2PThreads, 0, static_cast<cudaStream_t>(stream)>>>(a, static_cast<unsigned int*>(counter), gather);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

// offset of the gradient buffer inside a region (the Python side builds its flat_g tensor on it)
extern "C" size_t nwb_p2p_grad_offset(void) { return sizeof(P2PSignal); }
