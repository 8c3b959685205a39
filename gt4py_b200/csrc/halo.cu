// halo.cu -- halo refresh between the GPUs of one box over NVLink peer memory (new: the reference has no
// distributed layer).
//
// One process per GPU.  Every rank exports its field allocations with CUDA IPC; a rank refreshes its own
// halo by PULLING boxes straight out of the neighbours' memory with one kernel (loads on mapped peer
// pointers travel over NVLink / NVSwitch; no staging buffer, no NCCL call on the data path).
//
// Ordering between ranks uses step-numbered flags in peer-visible memory (system-scope release / acquire):
//   ready[r]   written by neighbour r into MY flag block: "r's field holds the data of step s"
//   done[r]    written by neighbour r into MY flag block: "r has finished pulling step s out of my field"
// b200_halo_refresh = { post ready(s) to every neighbour; wait ready(s) from every neighbour; copy the
// boxes; post done(s) to every neighbour }, all inside one kernel launched on the caller's stream, so
// "my data is ready" is simply stream order.  b200_halo_wait_done(s) is what a producer launches before it
// overwrites a field that neighbours pulled from in step s.  Each rank runs its kernels on its own GPU, so
// the spin-waits always make progress (never run two ranks of this protocol on ONE device).
#include <cstring>

#include "common.cuh"

namespace b200 {

struct HaloBox {
    const double *src;  // peer (or local) pointer of the box's first element
    double *dst;        // local pointer of the box's first element
    int64_t ss[3], ds[3];  // element strides
    int32_t n[3];          // extents (i, j, k)
    int32_t pad;           // element size in bytes (0 = 8)
};
static_assert(sizeof(HaloBox) == sizeof(b200_halo_box), "ABI struct mismatch");

constexpr int HALO_MAX_PEERS = 8;

struct HaloSync {
    // pointers into peer flag blocks where THIS rank posts, and local slots it waits on
    unsigned long long *post_ready[HALO_MAX_PEERS];
    unsigned long long *post_done[HALO_MAX_PEERS];
    const unsigned long long *wait_ready[HALO_MAX_PEERS];
    const unsigned long long *wait_done[HALO_MAX_PEERS];
    int npeers;
};
static_assert(sizeof(HaloSync) == sizeof(b200_halo_sync), "ABI struct mismatch");

__device__ __forceinline__ void post_flag(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long read_flag(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// grid = (blocks per box, nboxes).  *arrive (device counter, zero between launches) lets the last block
// post flags after every block finished copying.  mode:
//   0  plain box copy, no inter-rank protocol
//   1  PULL : post ready(s), wait neighbours' ready(s), copy (src = peer memory), post done(s)
//   2  PUSH : copy (dst = the neighbour's contiguous staging buffer), then post ready(s)
//   3  UNPACK: wait neighbours' ready(s), copy (local staging -> halo cells)
//   4  PUSH + UNPACK in ONE launch: boxes [0, nsplit) are pushed (the last of their blocks to finish posts
//      ready(s)), boxes [nsplit, nboxes) are unpacked by blocks that first wait for the neighbours' ready(s).
//      The push blocks have the lower block indices and the whole grid is far smaller than what a GPU keeps
//      resident, so the waiting blocks never keep a push block from running.
// PUSH / UNPACK is the fast path: boundary faces are only 2 elements wide along the contiguous axis, and
// 16-byte peer LOADS are bound by NVLink round-trip latency x outstanding requests, while the pushed stores
// are posted writes, coalesced because the staging buffer is contiguous.
// HaloBox.pad = element size in bytes (0 = 8): fields of every dtype (and, box by box, every data index) travel.
template <typename T>
__device__ __forceinline__ void copy_box_generic(const HaloBox &b)
{
    const T *src = reinterpret_cast<const T *>(b.src);
    T *dst = reinterpret_cast<T *>(b.dst);
    const int64_t total = (int64_t)b.n[0] * b.n[1] * b.n[2];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    auto off = [&](int64_t idx, const int64_t (&st)[3]) {
        const int i = (int)(idx % b.n[0]);
        const int64_t r = idx / b.n[0];
        return i * st[0] + (r % b.n[1]) * st[1] + (r / b.n[1]) * st[2];
    };
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += 2 * stride) {
        const int64_t idx2 = idx + stride;
        const T v0 = src[off(idx, b.ss)];
        T v1 = T();
        if (idx2 < total) v1 = src[off(idx2, b.ss)];
        dst[off(idx, b.ds)] = v0;
        if (idx2 < total) dst[off(idx2, b.ds)] = v1;
    }
}

__global__ void __launch_bounds__(256)
halo_refresh_kernel(const HaloBox *__restrict__ boxes, HaloSync sync, unsigned long long step,
                    unsigned int *arrive, int use_flags, int nsplit)
{
    int mode = use_flags;
    unsigned int push_blocks = gridDim.x * gridDim.y;
    if (mode == 4) {
        push_blocks = gridDim.x * (unsigned)nsplit;
        mode = (int)blockIdx.y < nsplit ? 2 : 3;
        // nothing to push (e.g. the rank at the open end of a one-sided halo): the neighbours still wait for this
        // rank's ready(s)
        if (nsplit == 0 && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0)
            for (int p = 0; p < sync.npeers; ++p) post_flag(sync.post_ready[p], step);
    }
    if (mode == 1 || mode == 3) {
        if (threadIdx.x == 0) {
            if (mode == 1 && blockIdx.x == 0 && blockIdx.y == 0)
                for (int p = 0; p < sync.npeers; ++p) post_flag(sync.post_ready[p], step);
            for (int p = 0; p < sync.npeers; ++p)
                while (read_flag(sync.wait_ready[p]) < step) __nanosleep(32);
        }
        __syncthreads();
    }
    const HaloBox b = boxes[blockIdx.y];
    const int64_t rows = (int64_t)b.n[1] * b.n[2];
    const int elem = b.pad ? b.pad : 8;
    if (elem == 8 && (b.n[0] & 1) == 0 && b.ss[0] == 1 && b.ds[0] == 1 &&
        ((reinterpret_cast<uintptr_t>(b.src) | reinterpret_cast<uintptr_t>(b.dst)) & 15) == 0 &&
        ((b.ss[1] | b.ss[2] | b.ds[1] | b.ds[2]) & 1) == 0) {
        // 128-bit path: pairs along the contiguous axis; two independent peer loads in flight per thread
        const int half = b.n[0] >> 1;
        const int64_t npairs = (int64_t)half * rows;
        const int64_t stride = (int64_t)gridDim.x * blockDim.x;
        auto src_of = [&](int64_t idx) {
            const int i2 = (int)(idx % half);
            const int64_t r = idx / half;
            return b.src + 2 * i2 + (r % b.n[1]) * b.ss[1] + (r / b.n[1]) * b.ss[2];
        };
        auto dst_of = [&](int64_t idx) {
            const int i2 = (int)(idx % half);
            const int64_t r = idx / half;
            return b.dst + 2 * i2 + (r % b.n[1]) * b.ds[1] + (r / b.n[1]) * b.ds[2];
        };
        for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < npairs; idx += 2 * stride) {
            const int64_t idx2 = idx + stride;
            const double2 v0 = *reinterpret_cast<const double2 *>(src_of(idx));
            double2 v1 = make_double2(0.0, 0.0);
            if (idx2 < npairs) v1 = *reinterpret_cast<const double2 *>(src_of(idx2));
            *reinterpret_cast<double2 *>(dst_of(idx)) = v0;
            if (idx2 < npairs) *reinterpret_cast<double2 *>(dst_of(idx2)) = v1;
        }
    } else if (elem == 8) {
        copy_box_generic<double>(b);
    } else if (elem == 4) {
        copy_box_generic<uint32_t>(b);
    } else if (elem == 2) {
        copy_box_generic<uint16_t>(b);
    } else {
        copy_box_generic<uint8_t>(b);
    }
    if (mode == 1 || mode == 2) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();  // this block's (peer) stores are visible before the flag
            if (atomicAdd(arrive, 1u) == push_blocks - 1) {
                *arrive = 0;  // ready for the next launch (stream order)
                __threadfence_system();
                for (int p = 0; p < sync.npeers; ++p)
                    post_flag(mode == 1 ? sync.post_done[p] : sync.post_ready[p], step);
            }
        }
    }
}

__global__ void halo_wait_done_kernel(HaloSync sync, unsigned long long step)
{
    if (threadIdx.x == 0)
        for (int p = 0; p < sync.npeers; ++p)
            while (read_flag(sync.wait_done[p]) < step) __nanosleep(64);
}

}  // namespace b200

extern "C" {

int b200_ipc_get_mem_handle(void *dptr, unsigned char handle[64])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    B200_REQUIRE(dptr && handle, "ipc_get_mem_handle: null argument");
    cudaIpcMemHandle_t h;
    B200_CUDA_CHECK(cudaIpcGetMemHandle(&h, dptr));
    memcpy(handle, &h, 64);
    return B200_OK;
}

int b200_ipc_open_mem_handle(const unsigned char handle[64], void **dptr)
{
    B200_REQUIRE(dptr && handle, "ipc_open_mem_handle: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    B200_CUDA_CHECK(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200_OK;
}

int b200_ipc_close_mem_handle(void *dptr)
{
    B200_CUDA_CHECK(cudaIpcCloseMemHandle(dptr));
    return B200_OK;
}

int b200_halo_refresh(const b200_halo_box *boxes_dev, int nboxes, int blocks_per_box,
                      const b200_halo_sync *sync, uint64_t step, unsigned int *arrive_dev, int use_flags,
                      void *stream)
{
    using namespace b200;
    // modes 0-3; mode 4 + 8 * nsplit: push boxes [0, nsplit) and unpack boxes [nsplit, nboxes) in one launch
    const int nsplit = use_flags >= 4 ? (use_flags - 4) / 8 : 0;
    if (use_flags >= 4) {
        B200_REQUIRE((use_flags - 4) % 8 == 0 && nsplit <= nboxes, "halo_refresh: bad fused-mode encoding %d", use_flags);
        use_flags = 4;
    }
    B200_REQUIRE(nboxes >= 0 && blocks_per_box > 0, "halo_refresh: bad launch shape");
    B200_REQUIRE(use_flags >= 0 && use_flags <= 4, "halo_refresh: unknown mode %d", use_flags);
    if (nboxes == 0 && !use_flags) return B200_OK;
    B200_REQUIRE(nboxes == 0 || boxes_dev, "halo_refresh: boxes is null");
    HaloSync s;
    memset(&s, 0, sizeof(s));
    if (use_flags) {
        B200_REQUIRE(sync && arrive_dev, "halo_refresh: sync block required with flags");
        B200_REQUIRE(sync->npeers >= 0 && sync->npeers <= HALO_MAX_PEERS, "halo_refresh: too many peers");
        memcpy(&s, sync, sizeof(s));
    }
    // with flags but no boxes (a rank without neighbours in this pattern) still post / wait
    dim3 grid(blocks_per_box, nboxes > 0 ? nboxes : 1);
    if (nboxes == 0) {
        static HaloBox *dummy = nullptr;
        if (!dummy) {
            B200_CUDA_CHECK(cudaMalloc(&dummy, sizeof(HaloBox)));
            B200_CUDA_CHECK(cudaMemset(dummy, 0, sizeof(HaloBox)));
        }
        boxes_dev = reinterpret_cast<const b200_halo_box *>(dummy);
        grid = dim3(1, 1);
    }
    halo_refresh_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const HaloBox *>(boxes_dev), s,
                                                                 (unsigned long long)step, arrive_dev, use_flags, nsplit);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}

int b200_halo_wait_done(const b200_halo_sync *sync, uint64_t step, void *stream)
{
    using namespace b200;
    B200_REQUIRE(sync && sync->npeers >= 0 && sync->npeers <= HALO_MAX_PEERS, "halo_wait_done: bad sync block");
    if (sync->npeers == 0) return B200_OK;
    HaloSync s;
    memcpy(&s, sync, sizeof(s));
    halo_wait_done_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(s, (unsigned long long)step);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}

}  // extern "C"
