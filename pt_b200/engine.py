"""ctypes front-end of the C ABI in include/ptgpu.h (pt_b200/lib/libptgpu.so).

Plumbing only: every method is one C-ABI call on host buffers.  There is no
Python or CPU implementation behind it -- if the CUDA library is missing or no
device is present, construction fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PTGPU_LIB") or os.path.join(_HERE, "lib", "libptgpu.so")

PTGPU_OK = 0
PTGPU_ERR_CUDA = -1
PTGPU_ERR_ARG = -2
PTGPU_ERR_STATE = -3
PTGPU_ERR_NOT_FINITE = -4

# == ptgpu_operation_t == pll_operation_t (libpll pll.h:249-259)
OP_DTYPE = np.dtype([
    ("parent_clv_index", np.uint32), ("parent_scaler_index", np.int32),
    ("child1_clv_index", np.uint32), ("child1_matrix_index", np.uint32),
    ("child1_scaler_index", np.int32),
    ("child2_clv_index", np.uint32), ("child2_matrix_index", np.uint32),
    ("child2_scaler_index", np.int32)])
CAND_DTYPE = np.dtype([("p", OP_DTYPE), ("q", OP_DTYPE), ("length0", np.float64)])
assert OP_DTYPE.itemsize == 32 and CAND_DTYPE.itemsize == 72

# ptgpu_allreduce_fn: int (*)(void *ctx, double *device_buf, size_t count, void *cuda_stream)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)

_dp = C.POINTER(C.c_double)
_up = C.POINTER(C.c_uint)
_lib = None


class PtGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"ptgpu error {code}: {msg}")
        self.code = code


def load_library():
    """dlopen the engine; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the engine has no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.ptgpu_last_error.restype = C.c_char_p
    L.ptgpu_stream.restype = C.c_void_p
    L.ptgpu_stream.argtypes = [C.c_void_p]
    L.ptgpu_destroy.argtypes = [C.c_void_p]
    L.ptgpu_destroy.restype = None
    _lib = L
    return L


def _d(a):
    return a.ctypes.data_as(_dp)


def comm_init(engines):
    """ptgpu_comm_init: the Engines of ONE process become the site shards 0..n-1 of one
    alignment (same device or peer-accessible devices); call set_shard on each as well"""
    L = load_library()
    arr = (C.c_void_p * len(engines))(*[e.h for e in engines])
    rc = L.ptgpu_comm_init(arr, C.c_uint(len(engines)))
    if rc != PTGPU_OK:
        raise PtGpuError(rc, L.ptgpu_last_error().decode())


def gamma_rates_mean(alpha, cats):
    """ptgpu_gamma_rates_mean = pll_compute_gamma_cats(alpha, cats, PLL_GAMMA_RATES_MEAN):
    host-only, needs no device"""
    L = load_library()
    out = np.empty(cats)
    rc = L.ptgpu_gamma_rates_mean(C.c_double(alpha), C.c_uint(cats), _d(out))
    if rc != PTGPU_OK:
        raise PtGpuError(rc, L.ptgpu_last_error().decode())
    return out


def ops_array(rows):
    """list of 8-int rows (pll_operation_t field order) -> structured array"""
    a = np.asarray(rows, dtype=np.int64).reshape(-1, 8)
    out = np.zeros(len(a), dtype=OP_DTYPE)
    for k, name in enumerate(OP_DTYPE.names):
        out[name] = a[:, k]
    return out


def compress_site_patterns(seqs):
    """ptgpu_compress_site_patterns (= pll_compress_site_patterns for nucleotide data): the
    unique columns of equally long sequences and their multiplicities.  Host-only."""
    L = load_library()
    n = len(seqs[0])
    bufs = [C.create_string_buffer(s.encode() if isinstance(s, str) else bytes(s), n + 1) for s in seqs]
    arr = (C.c_char_p * len(seqs))(*[C.cast(b, C.c_char_p) for b in bufs])
    w = np.zeros(max(n, 1), dtype=np.uint32)
    ln = C.c_uint(n)
    rc = L.ptgpu_compress_site_patterns(arr, C.c_uint(len(seqs)), C.byref(ln), w.ctypes.data_as(_up))
    if rc != PTGPU_OK:
        raise PtGpuError(rc, L.ptgpu_last_error().decode())
    m = ln.value
    return [b.raw[:m].decode() for b in bufs], w[:m].copy()


class Engine:
    """One alignment resident on one B200 (a ptgpu_engine handle)."""

    def __init__(self, tips, clv_buffers, sites, rate_cats, prob_matrices, scale_buffers,
                 device=0):
        self.lib = load_library()
        self.h = C.c_void_p()
        self.tips, self.clv_buffers, self.sites, self.cats = tips, clv_buffers, sites, rate_cats
        self.prob_matrices, self.scale_buffers = prob_matrices, scale_buffers
        self._chk(self.lib.ptgpu_create(C.c_uint(tips), C.c_uint(clv_buffers), C.c_uint(sites),
                                        C.c_uint(rate_cats), C.c_uint(prob_matrices),
                                        C.c_uint(scale_buffers), C.c_int(device),
                                        C.byref(self.h)))

    def _chk(self, rc):
        if rc != PTGPU_OK:
            raise PtGpuError(rc, self.lib.ptgpu_last_error().decode())

    def close(self):
        if self.h:
            self.lib.ptgpu_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_pattern_weights(self, weights):
        """pll_set_pattern_weights; None restores all-ones"""
        if weights is None:
            self._chk(self.lib.ptgpu_set_pattern_weights(self.h, None))
            return
        w = np.ascontiguousarray(weights, dtype=np.uint32)
        if w.size != self.sites:
            raise ValueError("one weight per site pattern")
        self._chk(self.lib.ptgpu_set_pattern_weights(self.h, w.ctypes.data_as(_up)))

    # --- model / data -----------------------------------------------------
    def set_model(self, freqs, subst, rates):
        f = np.ascontiguousarray(freqs, dtype=np.float64)
        s = np.ascontiguousarray(subst, dtype=np.float64)
        r = np.ascontiguousarray(rates, dtype=np.float64)
        assert len(f) == 4 and len(s) == 6 and len(r) == self.cats
        self._chk(self.lib.ptgpu_set_model(self.h, _d(f), _d(s), _d(r)))

    def eigen(self):
        ev = np.empty(16); iv = np.empty(16); el = np.empty(4)
        self._chk(self.lib.ptgpu_get_eigen(self.h, _d(ev), _d(iv), _d(el)))
        return ev, iv, el

    def set_tip_states(self, tip, seq):
        b = seq.encode() if isinstance(seq, str) else bytes(seq)
        if len(b) != self.sites:
            raise ValueError("sequence length does not match the partition")
        self._chk(self.lib.ptgpu_set_tip_states(self.h, C.c_uint(tip), b))

    def set_tip_states_ptr(self, tip, host_ptr):
        """same call, `host_ptr` = address of `sites` characters (e.g. pinned memory)"""
        self._chk(self.lib.ptgpu_set_tip_states(self.h, C.c_uint(tip), C.c_void_p(host_ptr)))

    def set_tip_states_batch(self, tips, host_ptrs):
        """all listed tips in one call; host_ptrs[i] = address of tip tips[i]'s characters"""
        t = np.ascontiguousarray(tips, dtype=np.uint32)
        ptrs = (C.c_void_p * len(t))(*[int(p) for p in host_ptrs])
        self._chk(self.lib.ptgpu_set_tip_states_batch(self.h, C.c_uint(len(t)),
                                                      t.ctypes.data_as(_up), ptrs))

    # --- hot path -----------------------------------------------------------
    def update_prob_matrices(self, indices, lengths):
        i = np.ascontiguousarray(indices, dtype=np.uint32)
        l = np.ascontiguousarray(lengths, dtype=np.float64)
        self._chk(self.lib.ptgpu_update_prob_matrices(self.h, i.ctypes.data_as(_up), _d(l),
                                                      C.c_uint(len(i))))

    def update_partials(self, ops):
        o = np.ascontiguousarray(ops, dtype=OP_DTYPE)
        self._chk(self.lib.ptgpu_update_partials(self.h, o.ctypes.data_as(C.c_void_p),
                                                 C.c_uint(len(o))))

    def edge_loglikelihood(self, pclv, pscaler, cclv, cscaler, matrix, persite=False):
        out = C.c_double()
        ps = np.empty(self.sites) if persite else None
        self._chk(self.lib.ptgpu_compute_edge_loglikelihood(
            self.h, C.c_uint(pclv), C.c_int(pscaler), C.c_uint(cclv), C.c_int(cscaler),
            C.c_uint(matrix), None if ps is None else _d(ps), C.byref(out)))
        return (out.value, ps) if persite else out.value

    def update_sumtable(self, pclv, cclv, pscaler=-1, cscaler=-1):
        self._chk(self.lib.ptgpu_update_sumtable(self.h, C.c_uint(pclv), C.c_uint(cclv),
                                                 C.c_int(pscaler), C.c_int(cscaler)))

    def derivatives(self, length):
        d1 = C.c_double(); d2 = C.c_double()
        self._chk(self.lib.ptgpu_compute_likelihood_derivatives(
            self.h, C.c_double(length), C.byref(d1), C.byref(d2)))
        return d1.value, d2.value

    def optimize_branch(self, pclv, pscaler, cclv, cscaler, matrix, length0, max_iter=32,
                        eps=1e-6, map_lambda=0.0):
        out = C.c_double(); it = C.c_int()
        self._chk(self.lib.ptgpu_optimize_branch(
            self.h, C.c_uint(pclv), C.c_int(pscaler), C.c_uint(cclv), C.c_int(cscaler),
            C.c_uint(matrix), C.c_double(length0), C.c_uint(max_iter), C.c_double(eps),
            C.c_double(map_lambda), C.byref(out), C.byref(it)))
        return out.value, it.value

    def optimize_branch_bounded(self, pclv, pscaler, cclv, cscaler, guess, xmin=1e-4, xmax=100.0,
                                tol=1e-5, max_iter=30):
        """(length, status, iterations): pllmod's safeguarded Newton on one edge"""
        out = C.c_double(); st = C.c_int(); it = C.c_int()
        self._chk(self.lib.ptgpu_optimize_branch_bounded(
            self.h, C.c_uint(pclv), C.c_int(pscaler), C.c_uint(cclv), C.c_int(cscaler),
            C.c_double(guess), C.c_double(xmin), C.c_double(xmax), C.c_double(tol),
            C.c_uint(max_iter), C.byref(out), C.byref(st), C.byref(it)))
        return out.value, st.value, it.value

    def score_nni_batch(self, cands, max_iter=32, eps=1e-6, map_lambda=0.0):
        c = np.ascontiguousarray(cands, dtype=CAND_DTYPE)
        n = len(c)
        lens = np.empty(n); lnl = np.empty(n); st = np.zeros(n, dtype=np.int32)
        self._chk(self.lib.ptgpu_score_nni_batch(
            self.h, c.ctypes.data_as(C.c_void_p), C.c_uint(n), C.c_uint(max_iter),
            C.c_double(eps), C.c_double(map_lambda), _d(lens), _d(lnl),
            st.ctypes.data_as(C.POINTER(C.c_int))))
        return lens, lnl, st

    def optimize_local_batch(self, prog, bl_min=1e-4, bl_max=100.0, tolerance=1e-6, smoothings=32):
        """all NNI candidates of a tree through pllmod_opt_optimize_branch_lengths_local;
        `prog` is a pt_b200.local_blo.Program.  Returns (lengths [B][NL], lnl, status,
        smoothings)."""
        n = prog.count
        types = np.ascontiguousarray(prog.step_types, dtype=np.int32)
        steps = np.ascontiguousarray(prog.steps)
        se = np.ascontiguousarray(prog.smooth_edges); fe = np.ascontiguousarray(prog.final_edges)
        l0 = np.ascontiguousarray(prog.lengths0, dtype=np.float64)
        lens = np.empty_like(l0); lnl = np.empty(n)
        st = np.zeros(n, dtype=np.int32); sm = np.zeros(n, dtype=np.int32)
        ip = C.POINTER(C.c_int)
        self._chk(self.lib.ptgpu_optimize_local_batch(
            self.h, C.c_uint(n), C.c_uint(prog.n_private_clvs), C.c_uint(prog.n_lengths),
            C.c_uint(prog.n_prologue), C.c_uint(prog.n_body), types.ctypes.data_as(ip),
            steps.ctypes.data_as(C.c_void_p), se.ctypes.data_as(C.c_void_p),
            fe.ctypes.data_as(C.c_void_p), _d(l0), C.c_double(bl_min), C.c_double(bl_max),
            C.c_double(tolerance), C.c_uint(smoothings), _d(lens), _d(lnl),
            st.ctypes.data_as(ip), sm.ctypes.data_as(ip)))
        return lens, lnl, st, sm

    def optimize_all_branches(self, steps, n_prologue, n_body, n_bodies, edge, lengths0,
                              prior_edges=(), max_iter=32, eps=1e-6, map_lambda=0.0, max_passes=33):
        """Partition::OptimizeAllBranches as one persistent kernel; `steps` is a STEP_DTYPE
        array on the engine's own index spaces (see include/ptgpu.h).  Returns
        (lengths [prob_matrices], lnl, passes)."""
        st = np.ascontiguousarray(steps)
        ed = np.ascontiguousarray(edge)
        l0 = np.ascontiguousarray(lengths0, dtype=np.float64)
        pr = np.ascontiguousarray(prior_edges, dtype=np.uint32)
        out = np.empty_like(l0); lnl = C.c_double(); status = C.c_int(); passes = C.c_int()
        self._chk(self.lib.ptgpu_optimize_all_branches(
            self.h, C.c_uint(n_prologue), C.c_uint(n_body), C.c_uint(n_bodies),
            st.ctypes.data_as(C.c_void_p), ed.ctypes.data_as(C.c_void_p), _d(l0),
            pr.ctypes.data_as(_up), C.c_uint(len(pr)), C.c_uint(max_iter), C.c_double(eps),
            C.c_double(map_lambda), C.c_uint(max_passes), _d(out), C.byref(lnl), C.byref(status),
            C.byref(passes)))
        return out, lnl.value, passes.value

    # --- test hooks -----------------------------------------------------------
    def get_clv(self, idx):
        out = np.empty(self.sites * self.cats * 4)
        self._chk(self.lib.ptgpu_get_clv(self.h, C.c_uint(idx), _d(out)))
        return out

    def set_clv(self, idx, data):
        a = np.ascontiguousarray(data, dtype=np.float64)
        assert a.size == self.sites * self.cats * 4
        self._chk(self.lib.ptgpu_set_clv(self.h, C.c_uint(idx), _d(a)))

    def get_scaler(self, idx):
        out = np.empty(self.sites, dtype=np.uint32)
        self._chk(self.lib.ptgpu_get_scaler(self.h, C.c_int(idx), out.ctypes.data_as(_up)))
        return out

    def get_pmatrix(self, idx):
        out = np.empty(self.cats * 16)
        self._chk(self.lib.ptgpu_get_pmatrix(self.h, C.c_uint(idx), _d(out)))
        return out

    def set_pmatrix(self, idx, data):
        a = np.ascontiguousarray(data, dtype=np.float64)
        assert a.size == self.cats * 16
        self._chk(self.lib.ptgpu_set_pmatrix(self.h, C.c_uint(idx), _d(a)))

    def get_sumtable(self):
        out = np.empty(self.sites * self.cats * 4)
        self._chk(self.lib.ptgpu_get_sumtable(self.h, _d(out)))
        return out

    def batch_sumtable_doubles(self):
        """doubles per site of a candidate's sumtable in score_nni_batch (16, or 13 compact)"""
        self.lib.ptgpu_batch_sumtable_doubles.restype = C.c_uint
        return int(self.lib.ptgpu_batch_sumtable_doubles(self.h))

    def counters(self):
        out = (C.c_ulonglong * 6)()
        self._chk(self.lib.ptgpu_get_counters(self.h, out))
        return dict(launches=out[0], site_updates=out[1], newton_passes=out[2], candidates=out[3],
                    lnl_sweeps=out[4], final_sweeps=out[5])

    def stream(self):
        return self.lib.ptgpu_stream(self.h)

    # --- site shards ----------------------------------------------------------
    def set_shard(self, site_offset, global_sites):
        self._chk(self.lib.ptgpu_set_shard(self.h, C.c_ulonglong(site_offset),
                                           C.c_ulonglong(global_sites)))

    def set_allreduce(self, fn):
        """fn: ALLREDUCE_FN instance (kept alive by the caller) or None"""
        self._hook = fn
        self._chk(self.lib.ptgpu_set_allreduce(
            self.h, C.cast(fn, C.c_void_p) if fn is not None else None, None))

    # the engines' own exchange over peer memory (no hook, no NCCL call): see include/ptgpu.h
    def shard_export(self):
        """64-byte handle of this shard's mailbox, to be sent to the other ranks"""
        buf = (C.c_ubyte * 64)()
        self._chk(self.lib.ptgpu_shard_export(self.h, buf))
        return bytes(buf)

    def shard_connect(self, rank, handles):
        """handles[r] = rank r's shard_export() (own entry ignored)"""
        blob = b"".join(bytes(h) for h in handles)
        assert len(blob) == 64 * len(handles)
        arr = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        self._chk(self.lib.ptgpu_shard_connect(self.h, C.c_uint(rank), C.c_uint(len(handles)), arr))

    def shard_disconnect(self):
        self._chk(self.lib.ptgpu_shard_disconnect(self.h))

    def shard_status(self):
        """(shard count, sticky time-out flag, exchanges completed)"""
        n = C.c_int(); err = C.c_int(); ex = C.c_ulonglong()
        self._chk(self.lib.ptgpu_shard_status(self.h, C.byref(n), C.byref(err), C.byref(ex)))
        return n.value, err.value, ex.value

    def synchronize(self):
        self._chk(self.lib.ptgpu_synchronize(self.h))

    KERNEL_KINDS = ("pmatrix", "partials", "edge_lnl", "sumtable", "newton", "nni_prepare",
                    "tip_encode")

    def set_profiling(self, on=True):
        self._chk(self.lib.ptgpu_set_profiling(self.h, C.c_int(int(on))))

    def kernel_times(self, reset=True):
        """{kind: (total ms, launches)} measured with CUDA events on the engine's stream"""
        ms = (C.c_double * 7)()
        n = (C.c_ulonglong * 7)()
        self._chk(self.lib.ptgpu_get_kernel_times(self.h, ms, n, C.c_int(int(reset))))
        return {k: (ms[i], int(n[i])) for i, k in enumerate(self.KERNEL_KINDS)}
