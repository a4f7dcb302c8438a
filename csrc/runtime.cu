// Host runtime: error reporting, device info, peer-mapped workspace management (CUDA IPC).
//
// The workspace replaces the reference's mpi4py buffers: `msgBuffers[j]` / the pre-posted
// Irecv targets (naive.py:68-79) become flagged, double-buffered slots in rank 0's memory that
// every worker GPU stores into directly over NVLink; the per-round `Isend(beta)` loop
// (naive.py:97-98) becomes peer stores into each GPU's `beta[2][d]` buffer.
#include <stdarg.h>
#include <atomic>
#include <stdlib.h>

#include "workspace.cuh"

namespace eh {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const DevInfo& dev_info() {
    static DevInfo cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        static DevInfo none;
        return none;
    }
    DevInfo& di = cache[dev];
    if (di.device != dev) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) {
            di.sm_count = prop.multiProcessorCount;
            di.coop = prop.cooperativeLaunch;
            di.device = dev;
        }
    }
    return di;
}

static std::atomic<int> g_reserved_sms[64];

int worker_sms() {
    const DevInfo& di = dev_info();
    if (di.device < 0) return 0;
    const int left = di.sm_count - g_reserved_sms[di.device].load(std::memory_order_relaxed);
    return left < 1 ? 1 : left;
}

uint32_t next_launch_seq() {
    // 1 .. 2^30-1, wraps; atomic so that host threads driving different devices can share it
    static std::atomic<uint32_t> seq{0};
    return (seq.fetch_add(1u, std::memory_order_relaxed) % ((1u << 30) - 1u)) + 1u;
}

int launch_grid_sync(const void* kern, dim3 grid, dim3 block, void** args, size_t smem, cudaStream_t stream) {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("EH_LAUNCH");
        mode = (e && strcmp(e, "plain") == 0) ? 1 : 0;
    }
    cudaError_t err = mode ? cudaLaunchKernel(kern, grid, block, args, smem, stream)
                           : cudaLaunchCooperativeKernel(kern, grid, block, args, smem, stream);
    if (err != cudaSuccess) {
        set_error("kernel launch failed: %s", cudaGetErrorString(err));
        return (int)err;
    }
    return EH_OK;
}

static inline uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

__global__ void ws_error_read_kernel(const uint32_t* err, int* out) { *out = (int)*err; }

}  // namespace eh

extern "C" const char* eh_last_error(void) { return eh::g_err; }
extern "C" int eh_version(void) { return 100; }

extern "C" int eh_device_info(int* sm_count, int* grad_dense_ctas, int* decoder_ctas) {
    const eh::DevInfo& di = eh::dev_info();
    EH_REQUIRE(di.device >= 0, "no CUDA device");
    if (sm_count) *sm_count = di.sm_count;
    if (grad_dense_ctas) *grad_dense_ctas = di.sm_count;
    if (decoder_ctas) *decoder_ctas = di.sm_count;
    return EH_OK;
}

extern "C" int eh_set_reserved_sms(int n) {
    const eh::DevInfo& di = eh::dev_info();
    EH_REQUIRE(di.device >= 0, "no CUDA device");
    EH_REQUIRE(n >= 0 && n < di.sm_count, "reserved SMs (%d) must be in 0..%d", n, di.sm_count - 1);
    eh::g_reserved_sms[di.device].store(n, std::memory_order_relaxed);
    return EH_OK;
}

extern "C" int eh_worker_sms(void) { return eh::worker_sms(); }

// ------------------------------------------------------------------------------------------
// Clock alignment.  %globaltimer is a per-GPU counter; message completion times stamped on a
// worker GPU are compared with the decoder GPU's own clock, so the offset of every peer is
// measured once at set-up by a ping-pong through the peer-mapped control words: the root stores
// seq into the peer's CTL_PING, the peer answers with its clock + seq in the root's CTL_PONG;
// offset = t_peer - (t_send + t_recv) / 2 of the exchange with the smallest round trip.
// ------------------------------------------------------------------------------------------
namespace eh {
__global__ void clock_sync_root_kernel(CtlWord* peer_ctl, CtlWord* my_ctl, uint32_t base, int iters,
                                       long long* out_offset, long long* out_rtt, unsigned long long timeout_ns) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    long long best_rtt = -1, best_off = 0;
    for (int k = 1; k <= iters; ++k) {
        const uint32_t seq = base + (uint32_t)k;
        const uint64_t t0 = globaltimer();
        st_release_sys(&peer_ctl[CTL_PING].v, seq);
        bool ok = true;
        while (ld_acquire_sys(&my_ctl[CTL_PONG].v) != seq) {
            if (globaltimer() - t0 > timeout_ns) { ok = false; break; }
        }
        if (!ok) break;
        const uint64_t t1 = globaltimer();
        const uint64_t tp = ld_arrival_time(&my_ctl[CTL_PONG].v);
        const long long rtt = (long long)(t1 - t0);
        if (best_rtt < 0 || rtt < best_rtt) {
            best_rtt = rtt;
            best_off = (long long)tp - (long long)(t0 + (t1 - t0) / 2);
        }
    }
    *out_offset = best_off;
    *out_rtt = best_rtt;
}
__global__ void clock_sync_peer_kernel(CtlWord* my_ctl, CtlWord* root_ctl, uint32_t base, int iters,
                                       unsigned long long timeout_ns) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const uint64_t t_begin = globaltimer();
    for (int k = 1; k <= iters; ++k) {
        const uint32_t seq = base + (uint32_t)k;
        while (ld_acquire_sys(&my_ctl[CTL_PING].v) != seq) {
            if (globaltimer() - t_begin > timeout_ns) return;
        }
        st_arrival_time(&root_ctl[CTL_PONG].v);
        st_release_sys(&root_ctl[CTL_PONG].v, seq);
    }
}
}  // namespace eh

extern "C" int eh_clock_sync_root(const eh_ws_layout* lay, void* d_ws_local, void* d_ws_peer, uint32_t base, int iters,
                                  long long* d_offset_ns, long long* d_rtt_ns, void* stream) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws_local && d_ws_peer && d_offset_ns && d_rtt_ns && iters >= 1, "clock_sync_root: bad argument");
    clock_sync_root_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ws_ctl(d_ws_peer, *lay), ws_ctl(d_ws_local, *lay), base, iters,
                                                             d_offset_ns, d_rtt_ns, 20ull * 1000000000ull);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_clock_sync_peer(const eh_ws_layout* lay, void* d_ws_local, void* d_ws_root, uint32_t base, int iters,
                                  void* stream) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws_local && d_ws_root && iters >= 1, "clock_sync_peer: bad argument");
    clock_sync_peer_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ws_ctl(d_ws_local, *lay), ws_ctl(d_ws_root, *lay), base, iters,
                                                             60ull * 1000000000ull);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

// ------------------------------------------------------------------------------------------
// Bridge to GPUs outside this NVLink domain (another node): their messages and parameters travel
// as stream-ordered NCCL point-to-point transfers into / out of the same workspace slots, and
// these two one-thread kernels, enqueued behind the receive, publish what a peer store would have
// published: the arrival record of a received message (stamped with this GPU's clock) and the
// parameter flag of a received parameter vector.
// ------------------------------------------------------------------------------------------
namespace eh {
__global__ void bridge_arrival_kernel(uint32_t* flag, uint32_t stamp) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        st_arrival_time(flag);
        st_release_sys(flag, stamp);
    }
}
__global__ void bridge_beta_flag_kernel(uint32_t* flag, uint32_t stamp) {
    if (threadIdx.x == 0 && blockIdx.x == 0) st_release_sys(flag, stamp);
}
}  // namespace eh

extern "C" void* eh_ws_slot_ptr(const eh_ws_layout* lay, void* d_ws, int worker, int msg, int parity) {
    if (!lay || !d_ws || worker < 0 || worker >= (int)lay->n_workers || msg < 0 || msg >= (int)lay->n_msgs) return nullptr;
    return eh::ws_slot(d_ws, *lay, worker, msg, parity & 1);
}
extern "C" void* eh_ws_beta_ptr(const eh_ws_layout* lay, void* d_ws, int parity) {
    if (!lay || !d_ws) return nullptr;
    return eh::ws_beta(d_ws, *lay, parity & 1);
}
extern "C" int eh_bridge_arrival(const eh_ws_layout* lay, void* d_ws, int worker, int msg, uint32_t stamp, void* stream) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws, "bridge_arrival: null argument");
    EH_REQUIRE(worker >= 0 && worker < (int)lay->n_workers && msg >= 0 && msg < (int)lay->n_msgs,
               "bridge_arrival: slot (%d,%d) outside the workspace", worker, msg);
    bridge_arrival_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ws_flag(d_ws, *lay, worker, msg), stamp);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
extern "C" int eh_bridge_beta_flag(const eh_ws_layout* lay, void* d_ws, uint32_t stamp, void* stream) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws, "bridge_beta_flag: null argument");
    bridge_beta_flag_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ws_beta_flag(d_ws, *lay), stamp);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_ws_layout_make(int n_workers, int n_msgs, int d, eh_ws_layout* out) {
    using namespace eh;
    EH_REQUIRE(out, "ws_layout: null out");
    EH_REQUIRE(n_workers >= 1 && n_workers <= EH_MAX_WORKERS, "ws_layout: n_workers=%d out of range", n_workers);
    EH_REQUIRE(n_msgs >= 1 && n_msgs <= EH_MAX_MSGS, "ws_layout: n_msgs=%d out of range", n_msgs);
    EH_REQUIRE(d >= 1, "ws_layout: d must be positive");
    memset(out, 0, sizeof(*out));
    const uint32_t d_pad = (uint32_t)align_up((uint64_t)d, 32);      // 128-byte rows
    out->n_workers = n_workers;
    out->n_msgs = n_msgs;
    out->d_pad = d_pad;
    uint64_t off = 0;
    out->off_ctl = off;
    off += sizeof(CtlWord) * CTL_COUNT;
    out->off_beta = off = align_up(off, 256);
    off += 2ull * d_pad * sizeof(float);
    out->off_flags = off = align_up(off, 256);
    off += (uint64_t)n_workers * EH_MAX_MSGS * 128;
    out->off_slots = off = align_up(off, 256);
    off += (uint64_t)n_workers * n_msgs * 2ull * d_pad * sizeof(float);
    out->bytes = align_up(off, 256);
    return EH_OK;
}

extern "C" int eh_ws_create(const eh_ws_layout* lay, void** d_ws, unsigned char ipc_handle[64]) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws, "ws_create: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void* p = nullptr;
    EH_CUDA(cudaMalloc(&p, lay->bytes));
    EH_CUDA(cudaMemset(p, 0, lay->bytes));
    if (ipc_handle) {
        cudaIpcMemHandle_t h;
        cudaError_t e = cudaIpcGetMemHandle(&h, p);
        if (e != cudaSuccess) {
            // single-process use does not need the handle; report but keep the allocation
            memset(ipc_handle, 0, 64);
            (void)cudaGetLastError();
        } else {
            memcpy(ipc_handle, &h, 64);
        }
    }
    *d_ws = p;
    return EH_OK;
}

extern "C" int eh_ws_open(const unsigned char ipc_handle[64], void** d_ws_peer) {
    using namespace eh;
    EH_REQUIRE(ipc_handle && d_ws_peer, "ws_open: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle, 64);
    EH_CUDA(cudaIpcOpenMemHandle(d_ws_peer, h, cudaIpcMemLazyEnablePeerAccess));
    return EH_OK;
}

extern "C" int eh_ws_close(void* d_ws_peer) {
    if (d_ws_peer) EH_CUDA(cudaIpcCloseMemHandle(d_ws_peer));
    return EH_OK;
}

extern "C" int eh_ws_destroy(void* d_ws) {
    if (d_ws) EH_CUDA(cudaFree(d_ws));
    return EH_OK;
}

extern "C" int eh_ws_reset(const eh_ws_layout* lay, void* d_ws, void* stream) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws, "ws_reset: null argument");
    // control words + flags; parameter buffers and slots need no clearing (stamps guard them)
    EH_CUDA(cudaMemsetAsync((char*)d_ws + lay->off_ctl, 0, sizeof(CtlWord) * CTL_COUNT, (cudaStream_t)stream));
    EH_CUDA(cudaMemsetAsync((char*)d_ws + lay->off_flags, 0, (size_t)lay->n_workers * EH_MAX_MSGS * 128,
                            (cudaStream_t)stream));
    return EH_OK;
}

extern "C" int eh_ws_error(const eh_ws_layout* lay, void* d_ws, int* err, void* stream) {
    using namespace eh;
    EH_REQUIRE(lay && d_ws && err, "ws_error: null argument");
    uint32_t v = 0;
    EH_CUDA(cudaMemcpyAsync(&v, ws_error(d_ws, *lay), sizeof(v), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    EH_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    *err = (int)v;
    return EH_OK;
}
