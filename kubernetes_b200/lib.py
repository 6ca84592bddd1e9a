"""ctypes bindings: the C-ABI engine (libksched_b200.so, include/ksched_b200.h) and the host
mirror / harness (libksched_host.so, csrc/host/gpu_scheduler.cpp), both built in-tree by
`make -C kubernetes_b200/csrc`. Nothing here computes a placement; a missing library is an error.
"""
import ctypes
import json
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
LIB_DIR = os.environ.get("KSB_LIB_DIR") or os.path.join(PKG, "lib")  # KSB_LIB_DIR: an instrumented build (make OUT=... FINE_TIMING=n)
ENGINE_SO = os.path.join(LIB_DIR, "libksched_b200.so")
HOST_SO = os.path.join(LIB_DIR, "libksched_host.so")

# every symbol include/ksched_b200.h declares
CABI_SYMBOLS = [
    "ksb_config_default", "ksb_create", "ksb_destroy", "ksb_last_error", "ksb_node_upsert", "ksb_node_remove",
    "ksb_pod_add", "ksb_pod_remove", "ksb_selector_owner_upsert", "ksb_selector_owner_remove", "ksb_schedule",
    "ksb_schedule_batch", "ksb_node_name", "ksb_device_id", "ksb_fit_error_message", "ksb_get_last_node_index",
    "ksb_set_last_node_index", "ksb_failed_predicates", "ksb_scores", "ksb_resync_begin", "ksb_resync_end", "ksb_reason_name", "ksb_node_info", "ksb_flush", "ksb_deferred_errors", "ksb_policy_set", "ksb_preempt", "ksb_preempt_victim", "ksb_schedule_extended", "ksb_cache_set_ttl", "ksb_cache_assume_pod", "ksb_cache_finish_binding", "ksb_cache_forget_pod",
    "ksb_cache_add_pod", "ksb_cache_update_pod", "ksb_cache_remove_pod", "ksb_cache_is_assumed_pod", "ksb_cache_cleanup_assumed",
    "ksb_cache_skip_pod_update", "ksb_get_stats", "ksb_replay_last_batch", "ksb_replay_last_batch_noalloc",
    "ksb_shard_export", "ksb_shard_connect",
]


class KsbConfig(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "device", "w_least_requested", "w_balanced", "w_selector_spread", "w_node_affinity", "w_taint_toleration",
        "w_inter_pod_affinity", "w_prefer_avoid", "hard_pod_affinity_weight", "label_words", "shard_rank", "shard_count",
        "group_size", "pipeline")]


class KsbResult(ctypes.Structure):
    _fields_ = [("status", ctypes.c_int32), ("node_index", ctypes.c_int32), ("num_filtered", ctypes.c_int32),
                ("num_nodes", ctypes.c_int32), ("score", ctypes.c_int64), ("device_mask", ctypes.c_uint32 * 4)]


def build():
    """Compile both libraries for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-s", "-C", os.path.join(PKG, "csrc")])


_engine = None
_host = None


def engine():
    """The C-ABI library. Raises if it was not built: there is no fallback path."""
    global _engine
    if _engine is None:
        if not os.path.exists(ENGINE_SO):
            raise RuntimeError("%s is missing: run `make -C kubernetes_b200/csrc` (no CPU fallback exists)" % ENGINE_SO)
        lib = ctypes.CDLL(ENGINE_SO, mode=ctypes.RTLD_GLOBAL)
        lib.ksb_create.restype = ctypes.c_void_p
        lib.ksb_create.argtypes = [ctypes.POINTER(KsbConfig), ctypes.c_char_p, ctypes.c_size_t]
        lib.ksb_destroy.argtypes = [ctypes.c_void_p]
        lib.ksb_last_error.restype = ctypes.c_char_p
        lib.ksb_last_error.argtypes = [ctypes.c_void_p]
        lib.ksb_config_default.argtypes = [ctypes.POINTER(KsbConfig)]
        for name in ("ksb_node_upsert", "ksb_pod_add"):
            getattr(lib, name).argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t]
        lib.ksb_schedule.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(KsbResult)]
        lib.ksb_schedule_batch.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_uint64),
                                           ctypes.c_int32, ctypes.POINTER(KsbResult)]
        lib.ksb_get_last_node_index.restype = ctypes.c_uint64
        lib.ksb_get_last_node_index.argtypes = [ctypes.c_void_p]
        _engine = lib
    return _engine


def host():
    """The host mirror + JSON harness (links against the C-ABI library)."""
    global _host
    if _host is None:
        engine()
        if not os.path.exists(HOST_SO):
            raise RuntimeError("%s is missing: run `make -C kubernetes_b200/csrc`" % HOST_SO)
        lib = ctypes.CDLL(HOST_SO)
        lib.ksbh_session_create.restype = ctypes.c_void_p
        lib.ksbh_session_create.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
        lib.ksbh_session_destroy.argtypes = [ctypes.c_void_p]
        lib.ksbh_session_run.restype = ctypes.c_void_p
        lib.ksbh_session_run.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        lib.ksbh_free.argtypes = [ctypes.c_void_p]
        lib.ksbh_cache_ops.restype = ctypes.c_void_p
        lib.ksbh_cache_ops.argtypes = [ctypes.c_char_p]
        _host = lib
    return _host


def cache_ops(request):
    """Drive the scheduler-cache state machine (host mirror of schedulercache.Cache over a fresh engine) with a list of
    operations; same request and reply as the oracle's "cache_ops" op."""
    h = host()
    p = h.ksbh_cache_ops(json.dumps(request).encode())
    try:
        out = json.loads(ctypes.string_at(p).decode())
    finally:
        h.ksbh_free(p)
    if "error" in out:
        raise RuntimeError("engine: " + out["error"])
    return out


class Session:
    """A parsed cluster + pod stream, pre-serialised into the C-ABI's wire records."""

    def __init__(self, cluster=None, path=None):
        h = host()
        err = ctypes.c_void_p()
        if path is not None:
            self._s = h.ksbh_session_create(path.encode(), 1, ctypes.byref(err))
        else:
            self._s = h.ksbh_session_create(json.dumps(cluster).encode(), 0, ctypes.byref(err))
        if not self._s:
            msg = ctypes.string_at(err).decode() if err else "unknown error"
            raise RuntimeError("ksbh_session_create: " + msg)

    def run(self, **options):
        """Run the stream on a fresh engine (the scheduleOne loop). Returns the harness's JSON as a dict."""
        h = host()
        p = h.ksbh_session_run(self._s, json.dumps(options).encode())
        try:
            out = json.loads(ctypes.string_at(p).decode())
        finally:
            h.ksbh_free(p)
        if "error" in out and "results" not in out and "pods" not in out:
            raise RuntimeError("engine: " + out["error"])
        return out

    def record(self, kind, index):
        """The wire record the C++ encoder built for one object (kind: 0 node, 1 pod of the stream, 2 existing pod,
        3 selector owner), or None. Tests hold it against kubernetes_b200/wire.py."""
        h = host()
        h.ksbh_session_record.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p),
                                          ctypes.POINTER(ctypes.c_size_t)]
        data, n = ctypes.c_void_p(), ctypes.c_size_t()
        if h.ksbh_session_record(self._s, kind, index, ctypes.byref(data), ctypes.byref(n)) != 0:
            return None
        return ctypes.string_at(data, n.value)

    def close(self):
        if self._s:
            host().ksbh_session_destroy(self._s)
            self._s = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class OpenEngine:
    """A loaded engine kept open across calls (needed for sharded runs: load, exchange CUDA-IPC handles
    between the ranks, then run). Wraps the ksbh_engine_* harness entry points."""

    def __init__(self, session, **options):
        h = host()
        h.ksbh_engine_open.restype = ctypes.c_void_p
        h.ksbh_engine_open.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]
        h.ksbh_engine_run.restype = ctypes.c_void_p
        h.ksbh_engine_run.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        h.ksbh_engine_close.argtypes = [ctypes.c_void_p]
        h.ksbh_engine_shard_export.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t,
                                               ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_void_p)]
        h.ksbh_engine_shard_connect.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_uint64),
                                                ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
        self._session = session  # keep the session alive
        err = ctypes.c_void_p()
        self._e = h.ksbh_engine_open(session._s, json.dumps(options).encode(), ctypes.byref(err))
        if not self._e:
            raise RuntimeError("ksbh_engine_open: " + (ctypes.string_at(err).decode() if err else "unknown error"))

    def shard_export(self, max_batch_pods):
        h = host()
        buf = ctypes.create_string_buffer(4096)
        n = ctypes.c_size_t()
        err = ctypes.c_void_p()
        if h.ksbh_engine_shard_export(self._e, int(max_batch_pods), buf, 4096, ctypes.byref(n), ctypes.byref(err)) != 0:
            raise RuntimeError("ksb_shard_export: " + (ctypes.string_at(err).decode() if err else "?"))
        return buf.raw[:n.value]

    def shard_connect(self, blobs):
        h = host()
        offs = [0]
        for b in blobs:
            offs.append(offs[-1] + len(b))
        arr = (ctypes.c_uint64 * len(offs))(*offs)
        err = ctypes.c_void_p()
        if h.ksbh_engine_shard_connect(self._e, b"".join(blobs), arr, len(blobs), ctypes.byref(err)) != 0:
            raise RuntimeError("ksb_shard_connect: " + (ctypes.string_at(err).decode() if err else "?"))

    def run(self, **options):
        h = host()
        p = h.ksbh_engine_run(self._e, json.dumps(options).encode())
        try:
            out = json.loads(ctypes.string_at(p).decode())
        finally:
            h.ksbh_free(p)
        if "error" in out and "pods" not in out:
            raise RuntimeError("engine: " + out["error"])
        return out

    def close(self):
        if self._e:
            host().ksbh_engine_close(self._e)
            self._e = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
