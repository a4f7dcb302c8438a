"""ctypes binding of ``liberasurehead.so`` (C ABI declared in ``include/erasurehead.h``).

The shared library is built in-tree by ``csrc/build.sh`` (nvcc, ``-gencode
arch=compute_100a,code=sm_100a``) and lives next to this file so that it travels with a
snapshot of the repository.  There is deliberately no fallback: on a machine with a GPU a
missing library is an error, not a silent switch to PyTorch eager code.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
LIB_PATH = os.environ.get("EH_LIB") or os.path.join(HERE, "liberasurehead.so")   # EH_LIB: A/B a different build
CSRC = os.path.join(ROOT, "csrc")

EH_MAX_WORKERS = 64
EH_MAX_MSGS = 2
EH_MAX_PEERS = 16
EH_MAX_SEGS = 16

_lock = threading.Lock()
_lib = None


class EhError(RuntimeError):
    pass


class WsLayout(C.Structure):
    _fields_ = [("bytes", C.c_uint64), ("off_ctl", C.c_uint64), ("off_beta", C.c_uint64),
                ("off_flags", C.c_uint64), ("off_slots", C.c_uint64),
                ("n_workers", C.c_uint32), ("n_msgs", C.c_uint32), ("d_pad", C.c_uint32),
                ("reserved", C.c_uint32)]


class GradJob(C.Structure):
    _fields_ = [("d_X", C.c_void_p * EH_MAX_SEGS), ("d_y", C.c_void_p * EH_MAX_SEGS),
                ("coeff", C.c_float * EH_MAX_SEGS),
                ("d_beta", C.c_void_p), ("d_out", C.c_void_p), ("d_flag", C.c_void_p),
                ("d_beta_flag", C.c_void_p), ("d_err", C.c_void_p),
                ("n_segs", C.c_int32), ("rows_per_seg", C.c_int32), ("d", C.c_int32), ("ld", C.c_int32),
                ("model", C.c_int32), ("stamp", C.c_uint32), ("timeout_ns", C.c_uint64),
                ("x_bf16", C.c_int32), ("abandon", C.c_int32)]


class GradEllJob(C.Structure):
    _fields_ = [("d_idx", C.c_void_p * EH_MAX_SEGS), ("d_y", C.c_void_p * EH_MAX_SEGS),
                ("coeff", C.c_float * EH_MAX_SEGS),
                ("d_beta", C.c_void_p), ("d_out", C.c_void_p), ("d_flag", C.c_void_p),
                ("d_beta_flag", C.c_void_p), ("d_err", C.c_void_p),
                ("n_segs", C.c_int32), ("rows_per_seg", C.c_int32), ("d", C.c_int32), ("nnz", C.c_int32),
                ("model", C.c_int32), ("bias_col", C.c_int32), ("stamp", C.c_uint32), ("reserved", C.c_uint32),
                ("timeout_ns", C.c_uint64)]


class WorkerJob(C.Structure):
    _fields_ = [("worker", C.c_int32), ("msg", C.c_int32), ("sparse", C.c_int32), ("model", C.c_int32),
                ("n_segs", C.c_int32), ("rows_per_seg", C.c_int32), ("d", C.c_int32), ("ld_or_nnz", C.c_int32),
                ("bias_col", C.c_int32), ("classes", C.c_int32),
                ("d_X", C.c_void_p * EH_MAX_SEGS), ("d_y", C.c_void_p * EH_MAX_SEGS),
                ("coeff", C.c_float * EH_MAX_SEGS)]


class McJob(C.Structure):
    _fields_ = [("d_X", C.c_void_p * EH_MAX_SEGS), ("d_y", C.c_void_p * EH_MAX_SEGS),
                ("coeff", C.c_float * EH_MAX_SEGS),
                ("d_W", C.c_void_p), ("d_out", C.c_void_p), ("d_loss", C.c_void_p), ("d_flag", C.c_void_p),
                ("d_beta_flag", C.c_void_p), ("d_err", C.c_void_p),
                ("n_segs", C.c_int32), ("rows_per_seg", C.c_int32), ("d", C.c_int32), ("ld", C.c_int32),
                ("ldw", C.c_int32), ("C", C.c_int32), ("model", C.c_int32), ("stamp", C.c_uint32),
                ("timeout_ns", C.c_uint64), ("x_bf16", C.c_int32), ("reserved", C.c_int32)]


class DecoderDesc(C.Structure):
    _fields_ = [("n_workers", C.c_int32), ("n_msgs", C.c_int32), ("n_groups", C.c_int32), ("s", C.c_int32),
                ("need_private", C.c_int32), ("need_count", C.c_int32), ("need_groups", C.c_int32),
                ("decode", C.c_int32), ("d", C.c_int32), ("d_pad", C.c_int32),
                ("update_agd", C.c_int32), ("n_peers", C.c_int32),
                ("d_group_of", C.c_void_p), ("d_B", C.c_void_p), ("d_A", C.c_void_p), ("d_binom", C.c_void_p),
                ("d_delay_ns", C.c_void_p), ("d_eta", C.c_void_p),
                ("d_beta64", C.c_void_p), ("d_u64", C.c_void_p),
                ("d_ws_local", C.c_void_p), ("d_ws_peer", C.c_void_p * EH_MAX_PEERS),
                ("d_betaset", C.c_void_p), ("d_timeset", C.c_void_p), ("d_worker_timeset", C.c_void_p),
                ("d_weights", C.c_void_p), ("d_ncollected", C.c_void_p), ("d_tstart", C.c_void_p),
                ("d_scratch", C.c_void_p), ("d_beta_mc", C.c_void_p),
                ("alpha", C.c_double), ("n_rows", C.c_double), ("lr_scale", C.c_double),
                ("timeout_ns", C.c_uint64), ("lay", WsLayout),
                ("d_clock_off", C.c_void_p), ("d_done", C.c_void_p), ("dec_ctas", C.c_int32), ("reserved", C.c_int32)]


class RoundPlan(C.Structure):
    _fields_ = [("n_jobs", C.c_int32), ("abandon", C.c_int32), ("jobs", C.POINTER(WorkerJob)),
                ("d_ws_local", C.c_void_p), ("d_ws_root", C.c_void_p), ("d_scratch", C.c_void_p),
                ("d_beta_local", C.c_void_p), ("scratch_bytes", C.c_size_t), ("decoder", C.POINTER(DecoderDesc)),
                ("timeout_ns", C.c_uint64), ("lay", WsLayout),
                ("dec_stream", C.c_void_p), ("ev_fork", C.c_void_p), ("ev_join", C.c_void_p)]


def _source_hash() -> str:
    """SHA-1 over the native sources and the public header (what the binary was built from)."""
    import hashlib
    h = hashlib.sha1()
    files = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".sh")))
    files.append(os.path.join(ROOT, "include", "erasurehead.h"))
    for f in files:
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a into ``liberasurehead.so`` (no GPU needed).  The library is
    rebuilt when it is missing or when the sources differ from the ones it was built from -- judged by a content
    hash kept next to it, not by file times (a copied tree, e.g. on a GPU box, has arbitrary times)."""
    stamp = LIB_PATH + ".srchash"
    want = _source_hash()
    have = open(stamp).read().strip() if os.path.exists(stamp) else ""
    stale = force or not os.path.exists(LIB_PATH) or have != want
    if stale:
        proc = subprocess.run(["bash", os.path.join(CSRC, "build.sh")], capture_output=True, text=True)
        if verbose or proc.returncode:
            print(proc.stdout[-4000:], proc.stderr[-4000:])
        if proc.returncode:
            raise EhError("nvcc build of liberasurehead.so failed")
        with open(stamp, "w") as f:
            f.write(want + "\n")
    return LIB_PATH


def _declare(lib):
    vp, i32, i64, u64, sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_size_t
    P = C.POINTER
    lib.eh_last_error.restype = C.c_char_p
    lib.eh_version.restype = i32
    sigs = {
        "eh_device_info": (i32, [P(i32), P(i32), P(i32)]),
        "eh_ws_layout_make": (i32, [i32, i32, i32, P(WsLayout)]),
        "eh_ws_create": (i32, [P(WsLayout), P(vp), C.c_char_p]),
        "eh_ws_open": (i32, [C.c_char_p, P(vp)]),
        "eh_ws_close": (i32, [vp]),
        "eh_ws_destroy": (i32, [vp]),
        "eh_ws_reset": (i32, [P(WsLayout), vp, vp]),
        "eh_ws_error": (i32, [P(WsLayout), vp, P(i32), vp]),
        "eh_ws_slot_ptr": (vp, [P(WsLayout), vp, i32, i32, i32]),
        "eh_ws_beta_ptr": (vp, [P(WsLayout), vp, i32]),
        "eh_bridge_arrival": (i32, [P(WsLayout), vp, i32, i32, C.c_uint32, vp]),
        "eh_bridge_beta_flag": (i32, [P(WsLayout), vp, C.c_uint32, vp]),
        "eh_grad_dense_scratch_bytes": (sz, [i32]),
        "eh_grad_dense": (i32, [P(GradJob), vp, sz, vp]),
        "eh_grad_dense_multi": (i32, [P(GradJob), i32, vp, sz, vp]),
        "eh_grad_dense_noalloc_max_ld": (sz, []),
        "eh_grad_dense_noalloc": (i32, [P(GradJob), vp]),
        "eh_grad_mc_scratch_bytes": (sz, [i64, i32, i32]),
        "eh_grad_mc": (i32, [P(McJob), vp, sz, vp]),
        "eh_grad_ell_scratch_bytes": (sz, [i32]),
        "eh_grad_ell": (i32, [P(GradEllJob), vp, sz, vp]),
        "eh_grad_ell_multi": (i32, [P(GradEllJob), i32, vp, sz, vp]),
        "eh_grad_ell_mc_padded_classes": (i32, []),
        "eh_grad_ell_mc_scratch_bytes": (sz, [i32]),
        "eh_grad_ell_mc_multi": (i32, [P(GradEllJob), i32, i32, vp, sz, vp]),
        "eh_decoder_scratch_bytes": (sz, []),
        "eh_decoder_ctas": (i32, [i32]),
        "eh_set_reserved_sms": (i32, [i32]),
        "eh_worker_sms": (i32, []),
        "eh_clock_sync_root": (i32, [P(WsLayout), vp, vp, C.c_uint32, i32, vp, vp, vp]),
        "eh_clock_sync_peer": (i32, [P(WsLayout), vp, vp, C.c_uint32, i32, vp]),
        "eh_decoder_round": (i32, [P(DecoderDesc), i32, i32, vp]),
        "eh_decoder_publish": (i32, [P(DecoderDesc), i32, vp]),
        "eh_run_rounds": (i32, [P(RoundPlan), i32, i32, i32, vp]),
        "eh_run_rounds_graph": (i32, [P(RoundPlan), i32, i32, i32, vp, P(vp)]),
        "eh_graph_destroy": (i32, [vp]),
        "eh_margins_dense": (i32, [vp, i64, i32, i32, vp, i32, i32, vp, vp]),
        "eh_margins_ell": (i32, [vp, i64, i32, vp, i32, i32, vp, vp]),
        "eh_eval_tc": (i32, [vp, i64, i32, i32, vp, i32, i32, vp, i32, vp, vp, i32, vp]),
        "eh_loss_columns": (i32, [vp, vp, i64, i32, i32, vp, vp]),
        "eh_auc": (i32, [vp, vp, i64, vp, vp, sz, vp]),
        "eh_auc_scratch_bytes": (sz, [i64]),
        "eh_auc_batch": (i32, [vp, vp, i64, i32, vp, vp, sz, vp]),
        "eh_auc_batch_scratch_bytes": (sz, [i64, i32]),
        "eh_text_shape": (i32, [C.c_char_p, P(i64), P(i64)]),
        "eh_load_dense_text": (i32, [C.c_char_p, i64, i64, vp, i64, i32]),
        "eh_save_dense_text": (i32, [C.c_char_p, vp, i64, i64, i64]),
        "eh_format_repr": (i32, [C.c_double, C.c_char_p]),
        "eh_synth_gmm": (i32, [vp, vp, i64, i32, i32, u64, u64, i32, i32, vp]),
        "eh_synth_onehot": (i32, [vp, vp, i64, i32, i32, vp, u64, u64, i32, i32, vp]),
        "eh_onehot_assemble": (i32, [vp, vp, i64, i32, vp, vp]),
        "eh_pair_codes": (i32, [vp, i64, i32, vp, i32, vp, vp, vp]),
        "eh_code_cyclic": (i32, [vp, i32, i32, vp, vp]),
        "eh_code_decode_table": (i32, [vp, i32, i32, vp, i64, vp, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:          # optional entry points are declared by whoever adds them
            continue
        fn.restype = res
        fn.argtypes = args


def lib():
    """Load (building first if necessary) and return the shared library."""
    global _lib
    with _lock:
        if _lib is None:
            # build() is a no-op when the library is newer than every source and header, so a stale
            # binary is never loaded against newer structs; a box without nvcc keeps what it has
            try:
                build()
            except (EhError, OSError):
                if not os.path.exists(LIB_PATH):
                    raise
            _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
            _declare(_lib)
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().eh_last_error().decode(errors="replace")
        raise EhError(f"{what or 'liberasurehead'} failed (rc={rc}): {msg}")


def ptr(t) -> int:
    """Device pointer of a torch tensor (or 0 for None)."""
    return 0 if t is None else t.data_ptr()


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
