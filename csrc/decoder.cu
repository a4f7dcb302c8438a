// Decoder round kernel (rank 0): arrival polling + stop rule + decode + update + parameter push.
//
// Role in the reference: the master's per-iteration body -- Isend beta to every worker, the
// Waitany loop with the scheme-specific stop rule, the decode (sum / first arrival per FRC
// group / least-squares row on the cyclic code) and the GD or AGD update
// (approximate_coding.py:136-180, coded.py:133-166, partial_coded.py:170-205,
// naive.py:112-124).  Here it is ONE kernel:
//
//   phase 1  (CTA 0, warp 0) polls the arrival records that worker kernels on any GPU publish
//            with st.release.sys.  A record carries the worker GPU's clock at the moment the
//            message was complete; the message becomes ready at  completion + injected delay
//            (delays.py) -- the reference worker's compute, sleep, Isend order
//            (approximate_coding.py:194-207) -- on the decoder's clock (per-GPU offsets measured
//            at set-up, eh_clock_sync_*).  Arrivals are consumed in ready-time order, the stop
//            rule is evaluated after each one and the decode weights are derived (table lookup by
//            combinadic rank, or an in-kernel pivoted elimination);
//   phase 2  (all CTAs) forms g = sum_w a_w * slot_w in float64, applies the update to the
//            float64 master copy, appends the iterate to betaset and pushes the float32
//            parameters into every GPU's workspace over NVLink (peer stores);
//   phase 3  (last CTA) releases the parameter flag of the next round on every GPU and closes
//            the round's timers.
//
// The kernel is a small grid (desc.dec_ctas CTAs) meant for its own high-priority stream next to
// the gradient kernels of the same GPU, which leave that many SMs free (eh_set_reserved_sms): the
// master only waits while the workers compute, as rank 0 of the reference does
// (approximate_coding.py:144-158), so a round costs max(compute, stop-rule wait), not their sum.
//
// Stragglers are never waited for: once the stop rule fires the late messages of this round
// are simply never read (their stamps no longer match), so the reference's stale-buffer hazards
//  cannot occur.
#include "workspace.cuh"
#include "solve.cuh"

namespace eh {

constexpr int DEC_THREADS = 512;
constexpr uint64_t NEVER = ~0ull;

struct DecScratch {
    alignas(128) uint32_t barrier;
    alignas(128) uint32_t done;
    alignas(128) uint32_t go;           // stamp once weights are final; high bit = abort
    alignas(128) double a_coded[EH_MAX_WORKERS];
    double a_priv[EH_MAX_WORKERS];
    int32_t n_collected;
    int32_t status;
};

struct Entry {             // one (worker, message) pair tracked by a polling lane
    uint64_t ready;
    int state;             // 0 waiting, 1 seen, 2 consumed, 3 unused
};

__global__ void __launch_bounds__(DEC_THREADS, 1)
decoder_round_kernel(const eh_decoder_desc D, const int round, const int r) {
    __shared__ double M[EH_MAX_WORKERS][EH_MAX_WORKERS + 1];
    __shared__ double s_a[EH_MAX_WORKERS];
    __shared__ int s_S[EH_MAX_WORKERS];
    __shared__ double s_wt[EH_MAX_WORKERS * EH_MAX_MSGS];
    __shared__ const float* s_src[EH_MAX_WORKERS * EH_MAX_MSGS];
    __shared__ int s_nsrc;
    __shared__ int s_cnt[2 * EH_MAX_WORKERS / 32];
    __shared__ uint32_t s_go;

    DecScratch* sc = reinterpret_cast<DecScratch*>(D.d_scratch);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t stamp = (uint32_t)round + 1u;
    const int parity = round & 1;
    const int n = D.n_workers;
    void* ws0 = D.d_ws_local;
    uint32_t* err = ws_error(ws0, D.lay);
    const uint64_t t_start = *D.d_tstart;

    // ------------------------------------------------------------------ phase 1: arrivals
    if (blockIdx.x == 0 && warp == 0) {
        const uint64_t* delay = D.d_delay_ns + (size_t)r * n;
        Entry ent[2][EH_MAX_MSGS];                 // workers lane, lane+32 ; msgs
        uint64_t dl[2];
        int grp_of[2];                             // FRC group of workers lane, lane + 32 (read once: a global load
                                                   // per consumed arrival sat in the serial consumption loop)
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int w = lane + 32 * q;
            dl[q] = (w < n) ? delay[w] : NEVER;
            grp_of[q] = (w < n) ? D.d_group_of[w] : 0;
#pragma unroll
            for (int m = 0; m < EH_MAX_MSGS; ++m) {
                ent[q][m].ready = NEVER;
                ent[q][m].state = (w < n && m < D.n_msgs) ? 0 : 3;
            }
        }
        unsigned long long done_mask = 0, first_mask = 0, priv_mask = 0, cov_mask = 0;
        int cnt_priv = 0, cnt_coded = 0, cnt_groups = 0, n_order = 0;
        bool aborted = false;
        bool stop = (cnt_priv >= D.need_private) && (cnt_coded >= D.need_count || cnt_groups >= D.need_groups);
        uint32_t idle = 0;
        bool drain = false;     // a message was just consumed: others may be ready too, skip the flag sweep
        while (!stop) {
            const uint64_t now = globaltimer();
            // poll the flags of entries still waiting (a message seen later can never be ready earlier
            // than one that is ready now, so draining the ready set first keeps the arrival order)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
#pragma unroll
                for (int m = 0; m < EH_MAX_MSGS; ++m) {
                    if (!drain && ent[q][m].state == 0) {
                        const int w = lane + 32 * q;
                        const uint32_t* rec = ws_flag(ws0, D.lay, w, m);
                        const uint32_t f = ld_acquire_sys(rec);
                        if (f == stamp) {
                            ent[q][m].state = 1;
                            // completion time on this GPU's clock, clamped into [round start, now]
                            uint64_t done = ld_arrival_time(rec);
                            if (D.d_clock_off) done -= (uint64_t)D.d_clock_off[w];
                            if ((int64_t)(done - t_start) < 0) done = t_start;
                            if (done > now) done = now;
                            ent[q][m].ready = (dl[q] == NEVER) ? NEVER : done + dl[q];
                            if (D.d_done && m == EH_MSG_CODED)
                                D.d_done[(size_t)r * n + w] = (double)(done - t_start) * 1e-9;
                        }
                    }
                }
            }
            // earliest ready message; ties: private before coded, then worker id
            uint64_t best = NEVER;
            uint32_t bkey = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
#pragma unroll
                for (int m = 0; m < EH_MAX_MSGS; ++m) {
                    if (ent[q][m].state == 1 && ent[q][m].ready <= now) {
                        const uint32_t key = ((m == EH_MSG_PRIVATE ? 0u : 1u) << 16) | (uint32_t)(lane + 32 * q);
                        if (ent[q][m].ready < best || (ent[q][m].ready == best && key < bkey)) {
                            best = ent[q][m].ready;
                            bkey = key;
                        }
                    }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const uint64_t ob = __shfl_xor_sync(0xffffffffu, best, o);
                const uint32_t ok = __shfl_xor_sync(0xffffffffu, bkey, o);
                if (ob < best || (ob == best && ok < bkey)) { best = ob; bkey = ok; }
            }
            if (best != NEVER) {
                idle = 0;
                drain = true;
                const int w = (int)(bkey & 0xffffu);
                const int m = (bkey >> 16) ? EH_MSG_CODED : EH_MSG_PRIVATE;
#pragma unroll
                for (int q = 0; q < 2; ++q)
#pragma unroll
                    for (int mm = 0; mm < EH_MAX_MSGS; ++mm)
                        if (lane + 32 * q == w && mm == m) ent[q][mm].state = 2;
                if (lane == 0 && D.d_worker_timeset)
                    D.d_worker_timeset[(size_t)r * n + w] = (double)(now - t_start) * 1e-9;
                if (m == EH_MSG_PRIVATE) {
                    cnt_priv += 1;
                    priv_mask |= 1ull << w;
                } else {
                    cnt_coded += 1;
                    done_mask |= 1ull << w;
                    if (lane == 0) s_S[n_order] = w;
                    n_order += 1;
                    const int g = __shfl_sync(0xffffffffu, (w & 32) ? grp_of[1] : grp_of[0], w & 31);
                    if (!((cov_mask >> g) & 1ull)) {
                        cov_mask |= 1ull << g;
                        cnt_groups += 1;
                        first_mask |= 1ull << w;
                    }
                }
                stop = (cnt_priv >= D.need_private) &&
                       (cnt_coded >= D.need_count || cnt_groups >= D.need_groups);
                continue;
            }
            // nothing ready: watchdog + back-off
            if (drain) { drain = false; continue; }
            const uint64_t limit = D.timeout_ns ? D.timeout_ns : 60ull * 1000000000ull;
            uint32_t e = 0;
            if (lane == 0) {
                e = ld_acquire_sys(err);
                if (e == 0 && now - t_start > limit) {
                    atomicCAS(err, 0u, (uint32_t)(-EH_ETIMEOUT));
                    e = (uint32_t)(-EH_ETIMEOUT);
                }
            }
            e = __shfl_sync(0xffffffffu, e, 0);
            if (e != 0) { aborted = true; break; }
            if (++idle > 8) nanosleep(idle > 64 ? 200 : 40);
        }
        __syncwarp();
        // ---- decode weights
        if (!aborted) {
            if (D.decode == EH_DECODE_LSTSQ) {
                const int k = D.need_count < n_order ? D.need_count : n_order;   // first n-s arrivals
                unsigned long long sel = 0;
                for (int c = 0; c < k; ++c) sel |= 1ull << s_S[c];
                for (int w = lane; w < n; w += 32) sc->a_coded[w] = 0.0;
                __syncwarp();
                if (D.d_A && D.d_binom && k == n - D.s) {
                    const long long rk = combinadic_rank(sel, n, D.d_binom);
                    for (int w = lane; w < n; w += 32) sc->a_coded[w] = D.d_A[(size_t)rk * n + w];
                } else {
                    solve_decode_row(D.d_B, n, s_S, k, M, s_a, lane);
                    for (int c = lane; c < k; c += 32) sc->a_coded[s_S[c]] = s_a[c];
                }
                // messages beyond the first n-s keep their arrival time but get weight 0
            } else {
                const unsigned long long use = (D.decode == EH_DECODE_FIRST_PER_GROUP) ? first_mask : done_mask;
                for (int w = lane; w < n; w += 32) sc->a_coded[w] = ((use >> w) & 1ull) ? 1.0 : 0.0;
            }
            for (int w = lane; w < n; w += 32) {
                sc->a_priv[w] = ((priv_mask >> w) & 1ull) ? 1.0 : 0.0;
                const bool collected = (done_mask >> w) & 1ull;
                if (D.d_worker_timeset && !collected) D.d_worker_timeset[(size_t)r * n + w] = -1.0;
            }
            __syncwarp();
            for (int w = lane; w < n; w += 32)
                if (D.d_weights) D.d_weights[(size_t)r * n + w] = sc->a_coded[w];
            if (lane == 0) {
                sc->n_collected = cnt_coded;
                if (D.d_ncollected) D.d_ncollected[r] = cnt_coded;
            }
        }
        __syncwarp();
        if (lane == 0) {
            __threadfence();
            st_release_gpu(&sc->go, aborted ? (stamp | 0x80000000u) : stamp);
        }
    }

    // ------------------------------------------------------------------ wait for the weights
    if (tid == 0) {
        uint32_t v;
        while (((v = ld_acquire_gpu(&sc->go)) & 0x7fffffffu) != stamp) nanosleep(64);
        s_go = v;
    }
    __syncthreads();
    const bool aborted = (s_go & 0x80000000u) != 0;

    // ------------------------------------------------------------------ phase 2: decode + update + push
    if (!aborted) {
        // source list (coded messages in worker order, then private ones): 128 threads fetch one weight each and
        // compact with ballots -- a serial loop of n dependent L2 loads by one thread cost ~0.4 us per worker
        {
            const int e = tid;                                   // entry: 0..63 coded, 64..127 private
            double a = 0.0;
            if (e < EH_MAX_WORKERS) {
                if (e < n) a = __ldcg(&sc->a_coded[e]);
            } else if (e < 2 * EH_MAX_WORKERS && D.n_msgs > 1) {
                if (e - EH_MAX_WORKERS < n) a = __ldcg(&sc->a_priv[e - EH_MAX_WORKERS]);
            }
            const bool nz = (tid < 2 * EH_MAX_WORKERS) && a != 0.0;
            const unsigned m = __ballot_sync(0xffffffffu, nz);
            if (tid < 2 * EH_MAX_WORKERS && lane == 0) s_cnt[warp] = __popc(m);
            __syncthreads();
            if (tid < 2 * EH_MAX_WORKERS) {
                int base = 0;
                for (int q = 0; q < warp; ++q) base += s_cnt[q];
                if (nz) {
                    const int c = base + __popc(m & ((1u << lane) - 1u));
                    const bool priv = e >= EH_MAX_WORKERS;
                    s_src[c] = ws_slot(ws0, D.lay, priv ? e - EH_MAX_WORKERS : e, priv ? EH_MSG_PRIVATE : EH_MSG_CODED, parity);
                    s_wt[c] = a;
                }
                if (tid == 0) {
                    int tot = 0;
                    for (int q = 0; q < 2 * EH_MAX_WORKERS / 32; ++q) tot += s_cnt[q];
                    s_nsrc = tot;
                }
            }
        }
        __syncthreads();
        const int nsrc = s_nsrc;
        const double eta = D.d_eta[r];
        const double gm = eta * D.lr_scale / D.n_rows;
        const double decay = 2.0 * D.alpha * eta;
        const double theta = 2.0 / ((double)round + 2.0);
        const int d4 = D.d_pad >> 2;
        const int nxt = (round + 1) & 1;
        for (int c4 = blockIdx.x * DEC_THREADS + tid; c4 < d4; c4 += gridDim.x * DEC_THREADS) {
            double g[4] = {0.0, 0.0, 0.0, 0.0};
            int q = 0;
            for (; q + 8 <= nsrc; q += 8) {            // eight loads in flight, summed in source order
                float4 t[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) t[u] = __ldcg(reinterpret_cast<const float4*>(s_src[q + u]) + c4);
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const double a = s_wt[q + u];
                    g[0] += a * (double)t[u].x; g[1] += a * (double)t[u].y;
                    g[2] += a * (double)t[u].z; g[3] += a * (double)t[u].w;
                }
            }
            for (; q + 4 <= nsrc; q += 4) {
                float4 t[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) t[u] = __ldcg(reinterpret_cast<const float4*>(s_src[q + u]) + c4);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const double a = s_wt[q + u];
                    g[0] += a * (double)t[u].x; g[1] += a * (double)t[u].y;
                    g[2] += a * (double)t[u].z; g[3] += a * (double)t[u].w;
                }
            }
            for (; q < nsrc; ++q) {
                const float4 t = __ldcg(reinterpret_cast<const float4*>(s_src[q]) + c4);
                const double a = s_wt[q];
                g[0] += a * (double)t.x; g[1] += a * (double)t.y; g[2] += a * (double)t.z; g[3] += a * (double)t.w;
            }
            float bn[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int c = c4 * 4 + e;
                bn[e] = 0.f;
                if (c < D.d) {
                    const double b = D.d_beta64[c];
                    double bnew;
                    if (D.update_agd) {
                        const double u = D.d_u64[c];
                        const double yt = (1.0 - theta) * b + theta * u;
                        bnew = yt - gm * g[e] - decay * b;
                        D.d_u64[c] = b + (bnew - b) * (1.0 / theta);
                    } else {
                        bnew = (1.0 - decay) * b - gm * g[e];
                    }
                    D.d_beta64[c] = bnew;
                    if (D.d_betaset) D.d_betaset[(size_t)r * D.d + c] = bnew;
                    bn[e] = (float)bnew;
                }
            }
            const float4 out = make_float4(bn[0], bn[1], bn[2], bn[3]);
            if (D.d_beta_mc) {
                multimem_st_v4(D.d_beta_mc + (size_t)nxt * D.d_pad + (size_t)c4 * 4, out);      // NVLS: fan-out in the switch
            } else {
                for (int pe = 0; pe < D.n_peers; ++pe)
                    *(reinterpret_cast<float4*>(ws_beta(D.d_ws_peer[pe], D.lay, nxt)) + c4) = out;
            }
        }
    }
    // ------------------------------------------------------------------ phase 3: publish
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        const uint32_t old = atom_add_acq_rel_gpu(&sc->done, 1u);
        if (old == gridDim.x - 1) {
            sc->done = 0;
            const uint64_t now = globaltimer();
            if (!aborted) {
                if (D.d_timeset) D.d_timeset[r] = (double)(now - t_start) * 1e-9;
                *D.d_tstart = now;
                __threadfence_system();
                for (int pe = 0; pe < D.n_peers; ++pe)
                    st_relaxed_sys(ws_beta_flag(D.d_ws_peer[pe], D.lay), stamp + 1u);   // one fence, n_peers stores
            }
        }
    }
}

// Initial parameter publication: beta64 -> every GPU's beta[round&1]; a second tiny kernel
// releases the flags so that the stores of the first are complete (kernel boundary) before.
__global__ void publish_push_kernel(const eh_decoder_desc D, const int round) {
    const int d4 = D.d_pad >> 2;
    const int par = round & 1;
    for (int c4 = blockIdx.x * blockDim.x + threadIdx.x; c4 < d4; c4 += gridDim.x * blockDim.x) {
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int c = c4 * 4 + e;
            v[e] = (c < D.d) ? (float)D.d_beta64[c] : 0.f;
        }
        const float4 out = make_float4(v[0], v[1], v[2], v[3]);
        if (D.d_beta_mc) {
            multimem_st_v4(D.d_beta_mc + (size_t)par * D.d_pad + (size_t)c4 * 4, out);
        } else {
            for (int pe = 0; pe < D.n_peers; ++pe)
                *(reinterpret_cast<float4*>(ws_beta(D.d_ws_peer[pe], D.lay, par)) + c4) = out;
        }
    }
    __threadfence_system();
}
__global__ void publish_flag_kernel(const eh_decoder_desc D, const int round) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        *D.d_tstart = globaltimer();
        __threadfence_system();
        for (int pe = 0; pe < D.n_peers; ++pe)
            st_release_sys(ws_beta_flag(D.d_ws_peer[pe], D.lay), (uint32_t)round + 1u);
    }
}

}  // namespace eh

extern "C" size_t eh_decoder_scratch_bytes(void) { return sizeof(eh::DecScratch); }

// CTAs (= SMs to reserve) a decoder for d_pad parameters wants: one float4 column per thread (the decode is a chain
// of dependent memory rounds per column, so columns per thread cost latency, not bandwidth), 1..16 CTAs, up to 32
// for parameter vectors beyond 128 K floats (the Amazon shape: rounds are bound by the decode of ~1 MB messages
// and the gradient kernels are small)
extern "C" int eh_decoder_ctas(int d_pad) {
    int c = (d_pad / 4 + eh::DEC_THREADS - 1) / eh::DEC_THREADS;
    const int cap = d_pad > 131072 ? 32 : 16;
    return c < 1 ? 1 : (c > cap ? cap : c);
}

static int check_desc(const eh_decoder_desc* D) {
    using namespace eh;
    EH_REQUIRE(D, "decoder: null descriptor");
    EH_REQUIRE(D->n_workers >= 1 && D->n_workers <= EH_MAX_WORKERS, "decoder: n_workers=%d out of range (1..%d)",
               D->n_workers, EH_MAX_WORKERS);
    EH_REQUIRE(D->n_msgs >= 1 && D->n_msgs <= EH_MAX_MSGS, "decoder: n_msgs out of range");
    EH_REQUIRE(D->n_peers >= 1 && D->n_peers <= EH_MAX_PEERS, "decoder: n_peers out of range");
    EH_REQUIRE(D->d > 0 && D->d_pad >= D->d && (D->d_pad % 4) == 0, "decoder: bad d/d_pad");
    EH_REQUIRE(D->d_beta64 && D->d_ws_local && D->d_scratch && D->d_tstart, "decoder: null state pointer");
    EH_REQUIRE(D->d_delay_ns && D->d_eta && D->d_group_of, "decoder: null table pointer");
    EH_REQUIRE(!D->update_agd || D->d_u64, "decoder: AGD needs d_u64");
    EH_REQUIRE(D->decode != EH_DECODE_LSTSQ || D->d_B, "decoder: lstsq decode needs B");
    for (int i = 0; i < D->n_peers; ++i) EH_REQUIRE(D->d_ws_peer[i], "decoder: null peer workspace %d", i);
    return EH_OK;
}

extern "C" int eh_decoder_round(const eh_decoder_desc* D, int round, int r, void* stream) {
    using namespace eh;
    int rc = check_desc(D);
    if (rc) return rc;
    const DevInfo& di = dev_info();
    int grid = (D->d_pad / 4 + DEC_THREADS - 1) / DEC_THREADS;
    const int cap = D->dec_ctas > 0 ? D->dec_ctas : di.sm_count;
    if (grid > cap) grid = cap;
    if (grid > di.sm_count) grid = di.sm_count;
    if (grid < 1) grid = 1;
    // CTAs > 0 only wait for CTA 0 (which is dispatched first) and nobody waits for them, so an ordinary
    // launch is enough; the grid is small so that it fits next to the gradient kernels of this GPU
    decoder_round_kernel<<<grid, DEC_THREADS, 0, (cudaStream_t)stream>>>(*D, round, r);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_decoder_publish(const eh_decoder_desc* D, int round, void* stream) {
    using namespace eh;
    int rc = check_desc(D);
    if (rc) return rc;
    int grid = (D->d_pad / 4 + 255) / 256;
    if (grid > 1024) grid = 1024;
    // load the round kernel now: with lazy module loading its first launch would otherwise wait for the gradient
    // kernels that are already running on the other stream, i.e. the first round would not overlap
    cudaFuncAttributes fa;
    EH_CUDA(cudaFuncGetAttributes(&fa, decoder_round_kernel));
    publish_push_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(*D, round);
    publish_flag_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*D, round);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
