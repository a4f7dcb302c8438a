// Native round loop (see eh_run_rounds in erasurehead.h).
//
// With plan->dec_stream set (root rank) the decoder kernels go to their own stream and overlap the
// gradient kernels of the same round on this GPU; the two streams are ordered against each other only
// through the stamped flags, plus one event fork at the start and one join at the end of the call.
//
// The host only *enqueues*: all cross-GPU ordering is carried by the stamped flags in the
// peer-mapped workspaces (worker kernels spin on the parameter flag, the decoder kernel spins on
// arrival flags), so `n_rounds` rounds are in flight on every GPU's stream with no host
// synchronisation -- the device-side analogue of the reference's pre-posted Irecv lists
// (approximate_coding.py:102-114) without their buffer-aliasing races.
#include "workspace.cuh"

extern "C" int eh_run_rounds(const eh_round_plan* plan, int first_round, int n_rounds, int r0, void* stream) {
    using namespace eh;
    EH_REQUIRE(plan && plan->d_ws_local && plan->d_ws_root && plan->d_scratch, "run_rounds: null pointer");
    EH_REQUIRE(plan->n_jobs == 0 || plan->jobs, "run_rounds: null job array");
    EH_REQUIRE(first_round >= 0 && n_rounds >= 0, "run_rounds: bad round range");
    const eh_ws_layout& lay = plan->lay;
    uint32_t* beta_flag = ws_beta_flag(plan->d_ws_local, lay);
    uint32_t* err = ws_error(plan->d_ws_local, lay);
    // decoder on its own stream: fork after everything already enqueued on `stream`, join at the end
    cudaStream_t dec_stream = (cudaStream_t)stream;
    const bool forked = plan->decoder && plan->dec_stream && plan->dec_stream != stream && n_rounds > 0;
    if (forked) {
        EH_REQUIRE(plan->ev_fork && plan->ev_join, "run_rounds: dec_stream needs ev_fork / ev_join");
        dec_stream = (cudaStream_t)plan->dec_stream;
        EH_CUDA(cudaEventRecord((cudaEvent_t)plan->ev_fork, (cudaStream_t)stream));
        EH_CUDA(cudaStreamWaitEvent(dec_stream, (cudaEvent_t)plan->ev_fork, 0));
    }
    for (int it = 0; it < n_rounds; ++it) {
        const int round = first_round + it;
        const int par = round & 1;
        const uint32_t stamp = (uint32_t)round + 1u;
        // dense single-vector messages of one shape are batched into one launch (<= 8 messages,
        // <= 32 partitions), and so are one-hot messages (<= 16 messages, <= 64 partitions); multi-column messages get their own kernels
        eh_grad_job batch[8];
        eh_grad_ell_job ebatch[16];
        int nb = 0, nb_segs = 0, ne = 0, ne_segs = 0, e_classes = 0;     // e_classes > 0: the one-hot batch is multi-column
        auto flush = [&]() -> int {
            int rc = EH_OK;
            if (nb > 0) rc = eh_grad_dense_multi(batch, nb, plan->d_scratch, plan->scratch_bytes, stream);
            nb = 0; nb_segs = 0;
            if (rc == EH_OK && ne > 0)
                rc = e_classes > 0 ? eh_grad_ell_mc_multi(ebatch, ne, e_classes, plan->d_scratch, plan->scratch_bytes, stream)
                                   : eh_grad_ell_multi(ebatch, ne, plan->d_scratch, plan->scratch_bytes, stream);
            ne = 0; ne_segs = 0; e_classes = 0;
            return rc;
        };
        for (int q = 0; q < plan->n_jobs; ++q) {
            const eh_worker_job& wj = plan->jobs[q];
            EH_REQUIRE(wj.worker >= 0 && wj.worker < (int)lay.n_workers && wj.msg >= 0 && wj.msg < (int)lay.n_msgs,
                       "run_rounds: job %d addresses slot (%d,%d) outside the workspace", q, wj.worker, wj.msg);
            float* out = ws_slot(plan->d_ws_root, lay, wj.worker, wj.msg, par);
            uint32_t* flag = ws_flag(plan->d_ws_root, lay, wj.worker, wj.msg);
            const float* beta = plan->d_beta_local ? plan->d_beta_local + (size_t)par * lay.d_pad
                                                   : ws_beta(plan->d_ws_local, lay, par);
            int rc = EH_OK;
            if (wj.sparse == 2 || wj.sparse == 4) {
                rc = flush();
                if (rc) return rc;
                eh_mc_job j;
                memset(&j, 0, sizeof(j));
                for (int i = 0; i < wj.n_segs && i < EH_MAX_SEGS; ++i) {
                    j.d_X[i] = (const float*)wj.d_X[i];
                    j.d_y[i] = wj.d_y[i];
                    j.coeff[i] = wj.coeff[i];
                }
                j.d_W = beta; j.d_out = out; j.d_flag = flag; j.d_beta_flag = beta_flag; j.d_err = err;
                j.n_segs = wj.n_segs; j.rows_per_seg = wj.rows_per_seg; j.d = wj.d; j.ld = wj.ld_or_nnz;
                j.ldw = (wj.d + 31) / 32 * 32; j.C = wj.classes; j.model = wj.model; j.stamp = stamp;
                j.timeout_ns = plan->timeout_ns;
                j.x_bf16 = (wj.sparse == 4);
                rc = eh_grad_mc(&j, plan->d_scratch, plan->scratch_bytes, stream);
            } else if (!wj.sparse || wj.sparse == 3) {
                eh_grad_job j;
                memset(&j, 0, sizeof(j));
                for (int i = 0; i < wj.n_segs && i < EH_MAX_SEGS; ++i) {
                    j.d_X[i] = wj.d_X[i];
                    j.d_y[i] = wj.d_y[i];
                    j.coeff[i] = wj.coeff[i];
                }
                j.d_beta = beta; j.d_out = out; j.d_flag = flag; j.d_beta_flag = beta_flag; j.d_err = err;
                j.n_segs = wj.n_segs; j.rows_per_seg = wj.rows_per_seg; j.d = wj.d; j.ld = wj.ld_or_nnz;
                j.model = wj.model; j.stamp = stamp; j.timeout_ns = plan->timeout_ns;
                j.x_bf16 = (wj.sparse == 3);
                j.abandon = plan->abandon;
                // batching pays for latency-bound messages; a bandwidth-bound message (> 16 MB of X) gets
                // its own launch so that it ships as soon as it is done instead of with the slowest one
                const size_t job_bytes = (size_t)j.n_segs * j.rows_per_seg * j.ld * (j.x_bf16 ? 2 : 4);
                const bool small = job_bytes <= ((size_t)16 << 20);
                const bool fits = small && nb < 8 && nb_segs + j.n_segs <= 32 &&
                                  (nb == 0 || (batch[0].d == j.d && batch[0].ld == j.ld &&
                                               batch[0].rows_per_seg == j.rows_per_seg && batch[0].model == j.model &&
                                               batch[0].x_bf16 == j.x_bf16 && (size_t)batch[0].n_segs * batch[0].rows_per_seg * batch[0].ld * sizeof(float) <=
                                                   ((size_t)16 << 20)));
                if (!fits) {
                    rc = flush();
                    if (rc) return rc;
                }
                batch[nb++] = j;
                nb_segs += j.n_segs;
            } else {
                eh_grad_ell_job j;
                memset(&j, 0, sizeof(j));
                for (int i = 0; i < wj.n_segs && i < EH_MAX_SEGS; ++i) {
                    j.d_idx[i] = (const int32_t*)wj.d_X[i];
                    j.d_y[i] = wj.d_y[i];
                    j.coeff[i] = wj.coeff[i];
                }
                j.d_beta = beta; j.d_out = out; j.d_flag = flag; j.d_beta_flag = beta_flag; j.d_err = err;
                j.n_segs = wj.n_segs; j.rows_per_seg = wj.rows_per_seg; j.d = wj.d; j.nnz = wj.ld_or_nnz;
                j.model = wj.model; j.bias_col = wj.bias_col; j.stamp = stamp; j.timeout_ns = plan->timeout_ns;
                const int cls = (wj.sparse == 5) ? wj.classes : 0;
                const bool fits = nb == 0 && ne < 16 && ne_segs + j.n_segs <= 64 && (ne == 0 || cls == e_classes) &&
                                  (ne == 0 || (ebatch[0].d == j.d && ebatch[0].nnz == j.nnz &&
                                               ebatch[0].rows_per_seg == j.rows_per_seg && ebatch[0].model == j.model &&
                                               ebatch[0].bias_col == j.bias_col));
                if (!fits) {
                    rc = flush();
                    if (rc) return rc;
                }
                ebatch[ne++] = j;
                ne_segs += j.n_segs;
                e_classes = cls;
            }
            if (rc) return rc;
        }
        {
            int rc = flush();
            if (rc) return rc;
        }
        if (plan->decoder) {
            int rc = eh_decoder_round(plan->decoder, round, r0 + it, (void*)dec_stream);
            if (rc) return rc;
        }
    }
    if (forked) {
        EH_CUDA(cudaEventRecord((cudaEvent_t)plan->ev_join, dec_stream));
        EH_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)plan->ev_join, 0));
    }
    return EH_OK;
}

extern "C" int eh_run_rounds_graph(const eh_round_plan* plan, int first_round, int n_rounds, int r0, void* stream,
                                   void** graph_exec) {
    using namespace eh;
    EH_REQUIRE(graph_exec, "run_rounds_graph: null output");
    cudaStream_t st = (cudaStream_t)stream;
    EH_REQUIRE(st != nullptr, "run_rounds_graph: capture needs a non-default stream");
    *graph_exec = nullptr;
    EH_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
    const int rc = eh_run_rounds(plan, first_round, n_rounds, r0, stream);
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(st, &graph);
    if (rc != EH_OK) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    EH_CUDA(e);
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    EH_CUDA(ei);
    const cudaError_t el = cudaGraphLaunch(exec, st);
    if (el != cudaSuccess) {
        cudaGraphExecDestroy(exec);
        EH_CUDA(el);
    }
    *graph_exec = (void*)exec;
    return EH_OK;
}

extern "C" int eh_graph_destroy(void* graph_exec) {
    if (graph_exec) EH_CUDA(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec));
    return EH_OK;
}
