"""Host-side binding of libdlco_b200.so (the C ABI in include/dlco_b200.h).

The reference (cbalint13/opencv-dlco) is a set of C++ command-line tools with no plugin API;
the functions here are the Python face of the C entry points that replace the tools' inline hot
loops (each docstring cites the reference loop).  torch is used only to hold device memory and
streams.  There is NO CPU fallback: importing works anywhere (so the symbol table can be
checked), but every compute call raises DlcoError unless the CUDA library initialises on an
sm_100 device.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdlco_b200.so")

c_void_p, c_int, c_int64, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
c_size_t, c_uint64, c_double, c_char_p = ctypes.c_size_t, ctypes.c_uint64, ctypes.c_double, ctypes.c_char_p
P = ctypes.POINTER

# name -> (restype, argtypes); mirrors include/dlco_b200.h one-to-one
SIGNATURES = {
    "dlco_init": (c_int, [c_int, P(c_void_p)]),
    "dlco_shutdown": (c_int, [c_void_p]),
    "dlco_last_error": (c_char_p, []),
    "dlco_device_name": (c_int, [c_void_p, c_char_p, c_int, P(c_int), P(c_int)]),
    "dlco_sync": (c_int, [c_void_p]),
    "dlco_set_stream": (c_int, [c_void_p, c_void_p]),
    "dlco_use_own_stream": (c_int, [c_void_p]),
    "dlco_malloc": (c_int, [c_void_p, c_size_t, P(c_void_p)]),
    "dlco_free": (c_int, [c_void_p, c_void_p]),
    "dlco_memcpy_h2d": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t]),
    "dlco_memcpy_d2h": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t]),
    "dlco_launch_count": (c_int64, [c_void_p]),
    "dlco_profile_enable": (c_int, [c_void_p, c_int]),
    "dlco_profile_read": (c_int, [c_void_p, c_int, P(c_double), P(c_int64)]),
    "dlco_bank_create": (c_int, [c_void_p, c_void_p, c_int, P(c_void_p)]),
    "dlco_bank_create_dev": (c_int, [c_void_p, c_void_p, c_int, P(c_void_p)]),
    "dlco_bank_destroy": (c_int, [c_void_p, c_void_p]),
    "dlco_ring_params": (c_int, [c_int, c_int, c_int, c_int, c_void_p, c_int, P(c_int)]),
    "dlco_gen_pr_filters": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, P(c_int)]),
    "dlco_bank_create_analytic": (c_int, [c_void_p, c_void_p, c_int, P(c_void_p)]),
    "dlco_bank_info": (c_int, [c_void_p, P(c_int), P(c_int), P(c_int), P(c_int64)]),
    "dlco_get_desc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p]),
    "dlco_fulldist_build": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int,
                                    c_float, c_int, c_void_p, c_int64]),
    "dlco_fulldist_build_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64,
                                         c_int64, c_int, c_float, c_int, c_void_p, c_int64]),
    "dlco_fulldist_build_host_async": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64,
                                               c_int64, c_int, c_float, c_int, c_void_p, c_int64, P(c_int64)]),
    "dlco_fulldist_host_wait": (c_int, [c_void_p, c_int64]),
    "dlco_pair_labels": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "dlco_uprjdist_build": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int,
                                    c_float, c_int, c_void_p, c_int64]),
    "dlco_hinge_pairsum": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_float, P(c_double)]),
    "dlco_pr_rda_run": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, ctypes.c_int32, c_void_p,
                                ctypes.c_int32, c_float, c_float, c_uint64, c_uint64, P(c_uint64),
                                c_void_p, c_void_p]),
    "dlco_pr_rda_run_grid": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, ctypes.c_int32, c_void_p,
                                     ctypes.c_int32, c_int, c_void_p, c_void_p, c_uint64, c_uint64,
                                     P(c_uint64), c_void_p, c_void_p]),
    "dlco_pr_validate": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int64,
                                 c_void_p, c_float, P(c_float), P(c_float)]),
    "dlco_pj_rda_grad": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                 c_int, c_void_p]),
    "dlco_pj_rda_grad_grid": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int, c_int,
                                      c_void_p, c_int64, c_void_p, c_void_p]),
    "dlco_pj_margins": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int, c_void_p]),
    "dlco_pj_hinge_weights": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "dlco_pj_wsyrk": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "dlco_pr_margins": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "dlco_pj_rda_update": (c_int, [c_void_p, c_int, c_uint64, c_int, c_float, c_float, c_void_p, c_void_p,
                                   c_void_p]),
    "dlco_pj_prox_apply": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, P(c_int)]),
    "dlco_pj_eig": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "dlco_pj_sample_batch": (c_int, [c_void_p, c_void_p, ctypes.c_int32, c_void_p, ctypes.c_int32, c_int, c_void_p,
                                     c_void_p, c_void_p]),
    "dlco_pj_rda_grad_grid_dev": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int, c_int,
                                          c_void_p, c_int64, c_void_p, c_void_p]),
    "dlco_pj_rda_update_grid": (c_int, [c_void_p, c_int, c_uint64, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p]),
    "dlco_pj_eig_batched": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "dlco_pj_prox_apply_grid": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                        c_void_p]),
    "dlco_pj_rda_step_grid": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, ctypes.c_int32, c_void_p,
                                      ctypes.c_int32, c_int, c_void_p, c_int, c_void_p, c_void_p, c_uint64,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dlco_pj_grid_check": (c_int, [c_void_p, c_int]),
    "dlco_pj_rda_step_grid_sign": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, ctypes.c_int32, c_void_p,
                                           ctypes.c_int32, c_int, c_void_p, c_int, c_void_p, c_void_p, c_uint64,
                                           c_void_p, c_void_p, c_void_p, c_void_p]),
    "dlco_pj_factor_grid": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "dlco_pj_psd_project": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "dlco_pj_validate": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int64,
                                 c_void_p, c_int, c_void_p, c_float, P(c_float), P(c_float)]),
    "dlco_pr_patchdist": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p]),
    "dlco_pj_patchdist": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p, c_int, c_void_p]),
    "dlco_make_current": (c_int, [c_void_p]),
    "dlco_comm_create_local": (c_int, [P(c_void_p), c_int, c_size_t, P(c_void_p)]),
    "dlco_comm_create": (c_int, [c_void_p, c_int, c_int, c_size_t, P(c_void_p)]),
    "dlco_comm_export": (c_int, [c_void_p, c_void_p]),
    "dlco_comm_import": (c_int, [c_void_p, c_int, c_void_p]),
    "dlco_comm_destroy": (c_int, [c_void_p]),
    "dlco_comm_allreduce_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t]),
    "dlco_comm_check": (c_int, [c_void_p]),
    "dlco_scatter_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int]),
    "dlco_gather_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "dlco_gemm3x_nt_f32": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p]),
    "dlco_roc_stats": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, P(c_float), P(c_double)]),
    "dlco_roc_stats_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, P(c_float), P(c_double)]),
    "dlco_roc_stats_static_max_n": (c_int64, []),
}


class DlcoError(RuntimeError):
    pass


def ring_params(patch_size=64, nr=32, nphi=5, nsigma=32):
    """"RingParams" of src/gen-poolregion.cpp:163-213 -> (n, 3) float64 (host arithmetic inside the library)."""
    import numpy as np
    lib = load_library()
    n = c_int()
    _check(lib.dlco_ring_params(patch_size, nr, nphi, nsigma, None, 0, ctypes.byref(n)))
    out = np.zeros((n.value, 3), np.float64)
    _check(lib.dlco_ring_params(patch_size, nr, nphi, nsigma, out.ctypes.data_as(c_void_p), n.value, ctypes.byref(n)))
    return out


_lib = None
MISSING = []


def load_library():
    """dlopen libdlco_b200.so and bind every entry point; raises if the library is missing
    (run `python -c 'import __graft_entry__ as g; g.build()'` first)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DlcoError("CUDA library %s is missing: build it first (no CPU fallback exists)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:           # tests/test_abi.py asserts MISSING stays empty
            MISSING.append(name)
            continue
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise DlcoError("libdlco_b200 error %d: %s" % (rc, load_library().dlco_last_error().decode()))


def _ptr(t):
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


class PeerComm:
    """dlco_comm for one-process-per-GPU launches (torch.distributed only carries the 64-byte IPC handles): the
    all-reduce itself is the library's peer-memory kernel, not a collective of the process group."""

    def __init__(self, ctx, max_count, group=None):
        import torch
        import torch.distributed as dist
        self.ctx, self.lib = ctx, load_library()
        self.rank, self.world = (dist.get_rank(group), dist.get_world_size(group)) if dist.is_initialized() else (0, 1)
        h = c_void_p()
        _check(self.lib.dlco_comm_create(ctx.handle, self.world, self.rank, max_count, ctypes.byref(h)))
        self.handle = h
        buf = ctypes.create_string_buffer(64)
        _check(self.lib.dlco_comm_export(h, buf))
        if self.world > 1:
            mine = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).to(ctx.device)
            allh = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(allh, mine, group=group)
            for p, t in enumerate(allh):
                if p != self.rank:
                    raw = bytes(t.cpu().numpy().tobytes())
                    _check(self.lib.dlco_comm_import(h, p, ctypes.create_string_buffer(raw, 64)))
            dist.barrier(group=group)

    def allreduce(self, t_in, t_out=None):
        t_out = t_in if t_out is None else t_out
        _check(self.lib.dlco_comm_allreduce_f32(self.handle, _ptr(t_in), _ptr(t_out), t_in.numel()))
        return t_out

    def check(self):
        _check(self.lib.dlco_comm_check(self.handle))

    def close(self):
        if self.handle:
            self.lib.dlco_comm_destroy(self.handle)
            self.handle = None


class Bank:
    def __init__(self, ctx, handle, n_filters):
        self.ctx, self.handle, self.n_filters = ctx, handle, n_filters

    def close(self):
        if self.handle:
            load_library().dlco_bank_destroy(self.ctx.handle, self.handle)
            self.handle = None

    def info(self):
        """-> dict(filters, rows (distinct), tiles (128 rows), k_chunks (non-empty 64-wide chunks))"""
        nf, nr, nt, kc = c_int(), c_int(), c_int(), c_int64()
        _check(load_library().dlco_bank_info(self.handle, ctypes.byref(nf), ctypes.byref(nr), ctypes.byref(nt),
                                            ctypes.byref(kc)))
        return {"filters": nf.value, "rows": nr.value, "tiles": nt.value, "k_chunks": kc.value}


class Context:
    """One per GPU (reference: cuda::setDevice(0), src/comp-fulldists.cpp:278)."""

    def __init__(self, device=0):
        import torch
        self.torch = torch
        self.lib = load_library()
        h = c_void_p()
        _check(self.lib.dlco_init(device, ctypes.byref(h)))
        self.handle = h
        self.device = torch.device("cuda", device)
        self.use_torch_stream()

    def use_torch_stream(self):
        s = self.torch.cuda.current_stream(self.device)
        _check(self.lib.dlco_set_stream(self.handle, c_void_p(s.cuda_stream)))

    def close(self):
        if self.handle:
            self.lib.dlco_shutdown(self.handle)
            self.handle = None

    def sync(self):
        _check(self.lib.dlco_sync(self.handle))

    def launch_count(self):
        return int(self.lib.dlco_launch_count(self.handle))

    def profile(self, on):
        _check(self.lib.dlco_profile_enable(self.handle, int(on)))

    def profile_read(self, which):
        ms, cnt = c_double(), c_int64()
        _check(self.lib.dlco_profile_read(self.handle, which, ctypes.byref(ms), ctypes.byref(cnt)))
        return ms.value, cnt.value

    def device_name(self):
        buf = ctypes.create_string_buffer(256)
        ma, mi = c_int(), c_int()
        _check(self.lib.dlco_device_name(self.handle, buf, 256, ctypes.byref(ma), ctypes.byref(mi)))
        return buf.value.decode(), ma.value, mi.value

    # ---- filter bank -------------------------------------------------------------------
    def bank(self, filters):
        """filters: (nF, 4096) float32, numpy (host) or torch cuda tensor."""
        torch = self.torch
        h = c_void_p()
        if isinstance(filters, torch.Tensor) and filters.is_cuda:
            f = filters.contiguous().float()
            _check(self.lib.dlco_bank_create_dev(self.handle, _ptr(f), f.shape[0], ctypes.byref(h)))
            torch.cuda.synchronize(self.device)
            return Bank(self, h, f.shape[0])
        import numpy as np
        f = np.ascontiguousarray(filters, dtype=np.float32).reshape(len(filters), -1)
        assert f.shape[1] == 4096
        _check(self.lib.dlco_bank_create(self.handle, f.ctypes.data_as(c_void_p), f.shape[0], ctypes.byref(h)))
        return Bank(self, h, f.shape[0])

    # ---- gen-poolregion: src/gen-poolregion.cpp:51-80,163-295 -----------------------------
    def gen_pr_filters(self, rings):
        """rings: (n, 3) float64 (r, phi, sigma) -> (PRFilters (8n, 4096) f32 cuda tensor, PRParams (8n, 3) f32
        numpy, number of filters the host libm had to regenerate)."""
        import numpy as np
        torch = self.torch
        rings = np.ascontiguousarray(rings, np.float64).reshape(-1, 3)
        out = torch.empty((len(rings) * 8, 4096), dtype=torch.float32, device=self.device)
        prp = np.zeros((len(rings) * 8, 3), np.float32)
        fixed = c_int()
        _check(self.lib.dlco_gen_pr_filters(self.handle, rings.ctypes.data_as(c_void_p), len(rings), _ptr(out),
                                            prp.ctypes.data_as(c_void_p), ctypes.byref(fixed)))
        return out, prp, fixed.value

    def bank_analytic(self, rings):
        """The bank of the given rings generated on the device (no host filters, no upload)."""
        import numpy as np
        rings = np.ascontiguousarray(rings, np.float64).reshape(-1, 3)
        h = c_void_p()
        _check(self.lib.dlco_bank_create_analytic(self.handle, rings.ctypes.data_as(c_void_p), len(rings),
                                                  ctypes.byref(h)))
        return Bank(self, h, len(rings) * 8)

    # ---- get_desc: src/vgg-desc.cpp:41-152 -----------------------------------------------
    def get_desc(self, patches, ids, n_bins=8, sigma=1.4, b_norm=True):
        torch = self.torch
        out = torch.empty((ids.numel(), 4096, 8), dtype=torch.float32, device=self.device)
        _check(self.lib.dlco_get_desc(self.handle, _ptr(patches), _ptr(ids), ids.numel(), n_bins,
                                      sigma, int(b_norm), _ptr(out)))
        return out

    # ---- comp-fulldists: src/comp-fulldists.cpp:294-356 ----------------------------------
    def fulldist_build(self, bank, patches, indices, lo=0, hi=None, n_bins=8, sigma=1.4, b_norm=True,
                       out=None):
        torch = self.torch
        hi = indices.shape[0] if hi is None else hi
        rings = bank.n_filters // 8
        if out is None:
            out = torch.empty((hi - lo, rings), dtype=torch.float32, device=self.device)
        _check(self.lib.dlco_fulldist_build(self.handle, bank.handle, _ptr(patches), _ptr(indices), lo, hi,
                                            n_bins, sigma, int(b_norm), _ptr(out), out.stride(0)))
        return out

    def fulldist_build_host(self, bank, patches, indices, lo=0, hi=None, n_bins=8, sigma=1.4,
                            b_norm=True, out=None):
        """patches / indices / out are HOST tensors (ideally pinned)."""
        torch = self.torch
        hi = indices.shape[0] if hi is None else hi
        rings = bank.n_filters // 8
        if out is None:
            out = torch.empty((hi - lo, rings), dtype=torch.float32).pin_memory()
        _check(self.lib.dlco_fulldist_build_host(self.handle, bank.handle, _ptr(patches), patches.shape[0],
                                                 _ptr(indices), lo, hi, n_bins, sigma, int(b_norm),
                                                 _ptr(out), out.stride(0)))
        return out

    def fulldist_build_host_async(self, bank, patches, indices, lo, hi, out, n_bins=8, sigma=1.4, b_norm=True):
        """Queues rows [lo, hi) into the HOST tensor `out` (hi - lo rows) without waiting; complete after sync() or
        fulldist_host_wait(ticket).  -> ticket"""
        t = c_int64()
        _check(self.lib.dlco_fulldist_build_host_async(self.handle, bank.handle, _ptr(patches), patches.shape[0],
                                                       _ptr(indices), lo, hi, n_bins, sigma, int(b_norm),
                                                       _ptr(out), out.stride(0), ctypes.byref(t)))
        return t.value

    def fulldist_host_wait(self, ticket):
        _check(self.lib.dlco_fulldist_host_wait(self.handle, ticket))

    def pair_labels(self, indices):
        out = self.torch.empty((indices.shape[0],), dtype=self.torch.uint8, device=self.device)
        _check(self.lib.dlco_pair_labels(self.handle, _ptr(indices), indices.shape[0], _ptr(out)))
        return out

    # ---- comp-uprjdists: src/comp-uprjdists.cpp:298-335 ----------------------------------
    def uprjdist_build(self, bank, patches, indices, lo=0, hi=None, n_bins=8, sigma=1.4, b_norm=True):
        torch = self.torch
        hi = indices.shape[0] if hi is None else hi
        out = torch.empty((hi - lo, bank.n_filters * 8), dtype=torch.float32, device=self.device)
        _check(self.lib.dlco_uprjdist_build(self.handle, bank.handle, _ptr(patches), _ptr(indices), lo, hi,
                                            n_bins, sigma, int(b_norm), _ptr(out), out.stride(0)))
        return out

    # ---- SubtractVectorsByRows + sum: src/kernelop-opencv.cu:49-80 -------------------------
    def hinge_pairsum(self, pos, neg, margin=1.0):
        s = c_double()
        _check(self.lib.dlco_hinge_pairsum(self.handle, _ptr(pos), pos.numel(), _ptr(neg), neg.numel(),
                                           margin, ctypes.byref(s)))
        return s.value

    # ---- pr-learn: src/pr-learn.cpp:303-358 ----------------------------------------------
    def pr_rda_run(self, dists, idx_pos, n_pos_trn, idx_neg, n_neg_trn, mu, gamma, t0, n_steps,
                   rng_state, w, dfavg):
        """Steps t0..t0+n_steps-1 of the l1/non-negative RDA; w, dfavg (F,) or (nCfg, F) float32
        cuda tensors updated in place; mu/gamma scalars or sequences (one per config).
        Returns the new cv::RNG state."""
        import numpy as np
        mus = np.atleast_1d(np.asarray(mu, np.float32))
        gammas = np.atleast_1d(np.asarray(gamma, np.float32))
        n_cfg = len(mus)
        assert len(gammas) == n_cfg and w.numel() == n_cfg * dists.shape[1] == dfavg.numel()
        st = c_uint64(rng_state)
        _check(self.lib.dlco_pr_rda_run_grid(
            self.handle, _ptr(dists), dists.stride(0), dists.shape[1], _ptr(idx_pos), n_pos_trn,
            _ptr(idx_neg), n_neg_trn, n_cfg, mus.ctypes.data_as(c_void_p), gammas.ctypes.data_as(c_void_p),
            t0, n_steps, ctypes.byref(st), _ptr(w), _ptr(dfavg)))
        return st.value

    def pr_validate(self, dists, pos_rows, neg_rows, w, mu):
        loss, regul = c_float(), c_float()
        _check(self.lib.dlco_pr_validate(self.handle, _ptr(dists), dists.stride(0), dists.shape[1],
                                         _ptr(pos_rows), pos_rows.numel(), _ptr(neg_rows), neg_rows.numel(),
                                         _ptr(w), mu, ctypes.byref(loss), ctypes.byref(regul)))
        return loss.value, regul.value

    # ---- pj-learn: src/pj-learn.cpp:311-527 ----------------------------------------------
    def pj_rda_grad(self, dists, rows_pos, rows_neg, W, rank, out=None):
        D = dists.shape[1]
        if out is None:
            out = self.torch.empty((D, D), dtype=self.torch.float32, device=self.device)
        _check(self.lib.dlco_pj_rda_grad(self.handle, _ptr(dists), dists.stride(0), D, _ptr(rows_pos),
                                         _ptr(rows_neg), rows_pos.numel(), _ptr(W), rank, _ptr(out)))
        return out

    def pj_rda_grad_grid(self, dists, rows_pos, rows_neg, W, ranks, out=None):
        """W: (nCfg, Rmax, D) cuda tensor, ranks: list of ints; -> (nCfg, D, D)."""
        D = dists.shape[1]
        n_cfg = len(ranks)
        if out is None:
            out = self.torch.empty((n_cfg, D, D), dtype=self.torch.float32, device=self.device)
        rk = (c_int * n_cfg)(*[int(r) for r in ranks])
        _check(self.lib.dlco_pj_rda_grad_grid(self.handle, _ptr(dists), dists.stride(0), D, _ptr(rows_pos),
                                              _ptr(rows_neg), rows_pos.numel(), n_cfg, _ptr(W),
                                              W.stride(0) if W is not None else 0, rk, _ptr(out)))
        return out

    # staged gradient (pair-sharded runs, shard.py)
    def pj_margins(self, dists, rows, W, rank):
        n = dists.shape[0] if rows is None else rows.numel()
        out = self.torch.empty((n,), dtype=self.torch.float32, device=self.device)
        if n == 0:
            return out
        _check(self.lib.dlco_pj_margins(self.handle, _ptr(dists), dists.stride(0), dists.shape[1], _ptr(rows), n,
                                        _ptr(W), rank, _ptr(out)))
        return out

    def pj_hinge_weights(self, margins):
        c = self.torch.empty_like(margins)
        _check(self.lib.dlco_pj_hinge_weights(self.handle, _ptr(margins), margins.numel() // 2, _ptr(c)))
        return c

    def pj_wsyrk(self, dists, rows, c, out=None):
        D = dists.shape[1]
        if out is None:
            out = self.torch.empty((D, D), dtype=self.torch.float32, device=self.device)
        _check(self.lib.dlco_pj_wsyrk(self.handle, _ptr(dists), dists.stride(0), D, _ptr(rows), rows.numel(),
                                      _ptr(c), _ptr(out)))
        return out

    def pr_margins(self, dists, rows, w):
        n = dists.shape[0] if rows is None else rows.numel()
        out = self.torch.empty((n,), dtype=self.torch.float32, device=self.device)
        if n == 0:
            return out
        _check(self.lib.dlco_pr_margins(self.handle, _ptr(dists), dists.stride(0), dists.shape[1], _ptr(rows), n,
                                        _ptr(w), _ptr(out)))
        return out

    def pj_rda_update(self, t, sz_batch, mu, gamma, d_loss, dfavg, A):
        _check(self.lib.dlco_pj_rda_update(self.handle, A.shape[0], t, sz_batch, mu, gamma, _ptr(d_loss),
                                           _ptr(dfavg), _ptr(A)))

    def pj_eig(self, A, evals=None, evecs=None):
        torch = self.torch
        D = A.shape[0]
        evals = torch.empty((D,), dtype=torch.float32, device=self.device) if evals is None else evals
        evecs = torch.empty((D, D), dtype=torch.float32, device=self.device) if evecs is None else evecs
        _check(self.lib.dlco_pj_eig(self.handle, D, _ptr(A), _ptr(evals), _ptr(evecs)))
        return evals, evecs

    def pj_prox_apply(self, evals, evecs, A, W):
        rank = c_int()
        _check(self.lib.dlco_pj_prox_apply(self.handle, A.shape[0], _ptr(evals), _ptr(evecs), _ptr(A), _ptr(W),
                                           ctypes.byref(rank)))
        return rank.value

    # ---- pj-learn sweep (nCfg configurations in lock step; state device resident) -------------
    def pj_sample_batch(self, idx_pos, n_pos_trn, idx_neg, n_neg_trn, sz_batch, rng_state, rows_pos, rows_neg):
        _check(self.lib.dlco_pj_sample_batch(self.handle, _ptr(idx_pos), n_pos_trn, _ptr(idx_neg), n_neg_trn,
                                             sz_batch, _ptr(rng_state), _ptr(rows_pos), _ptr(rows_neg)))

    def pj_rda_grad_grid_dev(self, dists, rows_pos, rows_neg, W, d_rank, out):
        n_cfg, D = W.shape[0], dists.shape[1]
        _check(self.lib.dlco_pj_rda_grad_grid_dev(self.handle, _ptr(dists), dists.stride(0), D, _ptr(rows_pos),
                                                  _ptr(rows_neg), rows_pos.numel(), n_cfg, _ptr(W), W.stride(0),
                                                  _ptr(d_rank), _ptr(out)))
        return out

    def pj_rda_update_grid(self, t, sz_batch, d_mu, d_gamma, d_loss, dfavg, A):
        _check(self.lib.dlco_pj_rda_update_grid(self.handle, A.shape[-1], t, sz_batch, A.shape[0], _ptr(d_mu),
                                                _ptr(d_gamma), _ptr(d_loss), _ptr(dfavg), _ptr(A)))

    def pj_eig_batched(self, A, evals=None, evecs=None):
        torch = self.torch
        n_cfg, D = A.shape[0], A.shape[-1]
        evals = torch.empty((n_cfg, D), dtype=torch.float32, device=self.device) if evals is None else evals
        evecs = torch.empty((n_cfg, D, D), dtype=torch.float32, device=self.device) if evecs is None else evecs
        _check(self.lib.dlco_pj_eig_batched(self.handle, D, n_cfg, _ptr(A), _ptr(evals), _ptr(evecs)))
        return evals, evecs

    def pj_prox_apply_grid(self, evals, evecs, A, W, d_rank):
        _check(self.lib.dlco_pj_prox_apply_grid(self.handle, A.shape[-1], A.shape[0], _ptr(evals), _ptr(evecs),
                                                _ptr(A), _ptr(W), W.stride(0), _ptr(d_rank)))

    def pj_rda_step_grid(self, dists, idx_pos, n_pos_trn, idx_neg, n_neg_trn, sz_batch, rng_state, d_mu, d_gamma,
                         t, W, A, dfavg, d_rank, work):
        _check(self.lib.dlco_pj_rda_step_grid(self.handle, _ptr(dists), dists.stride(0), dists.shape[1],
                                              _ptr(idx_pos), n_pos_trn, _ptr(idx_neg), n_neg_trn, sz_batch,
                                              _ptr(rng_state), W.shape[0], _ptr(d_mu), _ptr(d_gamma), t, _ptr(W),
                                              _ptr(A), _ptr(dfavg), _ptr(d_rank), _ptr(work)))

    def pj_rda_step_grid_sign(self, dists, idx_pos, n_pos_trn, idx_neg, n_neg_trn, sz_batch, rng_state, d_mu,
                              d_gamma, t, A, dfavg, d_rank, work):
        _check(self.lib.dlco_pj_rda_step_grid_sign(self.handle, _ptr(dists), dists.stride(0), dists.shape[1],
                                                   _ptr(idx_pos), n_pos_trn, _ptr(idx_neg), n_neg_trn, sz_batch,
                                                   _ptr(rng_state), A.shape[0], _ptr(d_mu), _ptr(d_gamma), t,
                                                   _ptr(A), _ptr(dfavg), _ptr(d_rank), _ptr(work)))

    def pj_factor_grid(self, A, d_rank, W, work):
        _check(self.lib.dlco_pj_factor_grid(self.handle, A.shape[-1], A.shape[0], _ptr(A), _ptr(d_rank), _ptr(W),
                                            W.stride(0), _ptr(work)))

    def pj_psd_project(self, A, want_sign=False):
        """In place: every symmetric matrix of A (n_cfg, D, D) becomes its PSD part; returns (ranks, sign)."""
        torch = self.torch
        n_cfg, D = A.shape[0], A.shape[-1]
        rank = torch.zeros((n_cfg,), dtype=torch.int32, device=self.device)
        S = torch.empty_like(A) if want_sign else None
        _check(self.lib.dlco_pj_psd_project(self.handle, D, n_cfg, _ptr(A), _ptr(rank), _ptr(S) if want_sign else None))
        return rank, S

    def pj_grid_check(self, n_cfg):
        _check(self.lib.dlco_pj_grid_check(self.handle, n_cfg))

    def pj_validate(self, dists, pos_rows, neg_rows, W, rank, A, mu):
        loss, regul = c_float(), c_float()
        _check(self.lib.dlco_pj_validate(self.handle, _ptr(dists), dists.stride(0), dists.shape[1],
                                         _ptr(pos_rows), pos_rows.numel(), _ptr(neg_rows), neg_rows.numel(),
                                         _ptr(W), rank, _ptr(A), mu, ctypes.byref(loss), ctypes.byref(regul)))
        return loss.value, regul.value

    # ---- FPR95 evaluation: src/misc.cpp:171-333 --------------------------------------------
    def pr_patchdist(self, dists, w):
        out = self.torch.empty((dists.shape[0],), dtype=self.torch.float32, device=self.device)
        _check(self.lib.dlco_pr_patchdist(self.handle, _ptr(dists), dists.shape[0], dists.stride(0),
                                          dists.shape[1], _ptr(w), _ptr(out)))
        return out

    def pj_patchdist(self, dists, W, rank):
        out = self.torch.empty((dists.shape[0],), dtype=self.torch.float32, device=self.device)
        _check(self.lib.dlco_pj_patchdist(self.handle, _ptr(dists), dists.shape[0], dists.stride(0),
                                          dists.shape[1], _ptr(W), rank, _ptr(out)))
        return out

    def roc_stats(self, patch_dist, labels):
        """-> (FPR95 float32, AUC float64)"""
        f, a = c_float(), c_double()
        _check(self.lib.dlco_roc_stats(self.handle, _ptr(patch_dist), _ptr(labels), patch_dist.numel(),
                                       ctypes.byref(f), ctypes.byref(a)))
        return f.value, a.value

    def roc_stats_static(self, patch_dist, labels):
        """roc_stats through dlco_roc_stats_static: scratch in a fixed __device__ array, nothing allocated; at most
        lib.dlco_roc_stats_static_max_n() scores.  Not in the reference (see include/dlco_b200.h)."""
        f, a = c_float(), c_double()
        _check(self.lib.dlco_roc_stats_static(self.handle, _ptr(patch_dist), _ptr(labels), patch_dist.numel(),
                                              ctypes.byref(f), ctypes.byref(a)))
        return f.value, a.value

    def gemm3x_nt(self, A, B):
        """A (M,K) . B (N,K)^T on tcgen05 with the 3xTF32 split (diagnostic entry point)."""
        out = self.torch.empty((A.shape[0], B.shape[0]), dtype=self.torch.float32, device=self.device)
        _check(self.lib.dlco_gemm3x_nt_f32(self.handle, _ptr(A.contiguous()), A.shape[0], _ptr(B.contiguous()),
                                           B.shape[0], A.shape[1], _ptr(out)))
        return out
