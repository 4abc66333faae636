"""Host-side mirror of the reference's pybind11 module (bindings/python/sparsesolvers/binding.cpp).

    Homotopy(A).solve(b, tolerance=10*eps, max_iterations=100) -> (x, HomotopyReport)    binding.cpp:79-111
    Irls(A).solve(b, ...)                                      -> (x, IrlsReport)
    version() -> [major, minor, patch]                                                   binding.cpp:42-46

A may be any 2-D float32/float64 array view (row subsets, column subsets, transposes: its strides are
passed through, binding.cpp:21-37); b must have the dictionary's dtype exactly (`noconvert`, :108).
`solve_batch` is additive: many signals against one dictionary in one call.
"""
import ctypes as C

import numpy as np

from ._lib import SolverError, check, lib

__all__ = ["Homotopy", "HomotopyGroup", "Irls", "HomotopyReport", "IrlsReport", "version", "SolverError",
           "reconstruct_signal", "norm_l1", "norm_l1_noalloc", "NOALLOC_SCRATCH_BYTES", "fp64_peak"]


def version():
    v = (C.c_int * 3)()
    lib().ssb200_version(v)
    return list(v)


class HomotopyReport:
    """include/ss/policies.h:25-32"""
    def __init__(self, iter=0, solution_error=0.0):
        self.iter = int(iter)
        self.solution_error = float(solution_error)

    def __repr__(self):
        return f"HomotopyReport(iter={self.iter}, solution_error={self.solution_error})"


class IrlsReport:
    """include/ss/policies.h:58-71"""
    def __init__(self, iter=0, solution_error=0.0, spd_failure=False):
        self.iter = int(iter)
        self.solution_error = float(solution_error)
        self.spd_failure = bool(spd_failure)

    def __repr__(self):
        return (f"IrlsReport(iter={self.iter}, solution_error={self.solution_error}, "
                f"spd_failure={self.spd_failure})")


_SFX = {np.dtype(np.float32): ("f32", C.c_float), np.dtype(np.float64): ("f64", C.c_double)}


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


def _is_cuda(a):
    """A torch CUDA tensor, or any other device array (CuPy, Numba, ...) exposing __cuda_array_interface__ / DLPack."""
    if type(a).__module__.split(".")[0] == "torch":
        return bool(getattr(a, "is_cuda", False))
    return hasattr(a, "__cuda_array_interface__")


def _as_torch(a):
    import torch
    return a if isinstance(a, torch.Tensor) else torch.from_dlpack(a) if hasattr(a, "__dlpack__") else torch.as_tensor(a, device="cuda")


class _Solver:
    _name = None

    def __init__(self, A, device=0):
        if _is_cuda(A):
            self._init_from_cuda(_as_torch(A))
            return
        A = np.asarray(A)
        if A.ndim != 2:
            raise ValueError("A must be 2-dimensional")
        if A.dtype not in _SFX:
            raise TypeError("A must be float32 or float64")          # binding.cpp:82-85 overloads
        self._sfx, self._ct = _SFX[A.dtype]
        self.dtype = A.dtype
        self.shape = A.shape
        self.device = device
        self._h = C.c_void_p()
        rs, cs = A.strides[0] // A.itemsize, A.strides[1] // A.itemsize
        fn = getattr(lib(), f"ssb200_{self._name}_create_{self._sfx}")
        check(fn(_ptr(A), C.c_size_t(A.shape[0]), C.c_size_t(A.shape[1]), C.c_ssize_t(rs), C.c_ssize_t(cs),
                 C.c_int(device), C.byref(self._h)))

    def _init_from_cuda(self, A):
        """Dictionary already on the device (a torch CUDA tensor or any DLPack / __cuda_array_interface__ producer): no host copy."""
        import torch
        if self._name != "homotopy":
            raise TypeError("a device-resident dictionary is supported by Homotopy only (the IRLS set-up factorises on the host)")
        if A.ndim != 2:
            raise ValueError("A must be 2-dimensional")
        dt = {torch.float32: np.dtype(np.float32), torch.float64: np.dtype(np.float64)}.get(A.dtype)
        if dt is None:
            raise TypeError("A must be float32 or float64")
        self._sfx, self._ct = _SFX[dt]
        self.dtype, self.shape, self.device = dt, tuple(A.shape), A.device.index
        self._h = C.c_void_p()
        fn = getattr(lib(), f"ssb200_{self._name}_create_dev_{self._sfx}")
        check(fn(C.c_void_p(A.data_ptr()), C.c_size_t(A.shape[0]), C.c_size_t(A.shape[1]), C.c_ssize_t(A.stride(0)),
                 C.c_ssize_t(A.stride(1)), C.c_int(self.device), C.byref(self._h)))

    def _check_cuda_signals(self, Y):
        import torch
        want = torch.float32 if self.dtype == np.float32 else torch.float64
        if Y.dtype != want:
            raise TypeError(f"signals must have dtype {want}")                         # py::arg().noconvert()
        if Y.ndim != 2 or Y.shape[1] != self.shape[0]:
            raise ValueError(f"signals must be (batch, {self.shape[0]})")
        if Y.device.index != self.device:
            raise ValueError(f"signals live on cuda:{Y.device.index}, the solver on cuda:{self.device}")
        return Y if Y.stride(1) == 1 and Y.stride(0) >= Y.shape[1] else Y.contiguous()

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            lib().ssb200_destroy(h)
            self._h = C.c_void_p()

    def _check_signal(self, b, ndim):
        if not isinstance(b, np.ndarray) or b.dtype != self.dtype:
            raise TypeError(f"signal must be a numpy array of dtype {self.dtype}")   # py::arg().noconvert()
        if b.ndim != ndim or b.shape[-1] != self.shape[0]:
            raise ValueError(f"signal must have {self.shape[0]} entries")

    def _defaults(self, tolerance, max_iterations):
        if tolerance is None:
            tolerance = 10 * np.finfo(self.dtype).eps                # binding.cpp:94
        return self._ct(tolerance), C.c_uint32(int(max_iterations))

    # instrumentation used by bench.py and the tests
    def counters(self):
        c = (C.c_uint64 * 4)()
        check(lib().ssb200_get_counters(self._h, c))
        return dict(launches=c[0], contractions=c[1], contraction_rows=c[2], iterations=c[3])

    def reset_counters(self):
        check(lib().ssb200_reset_counters(self._h))

    def set_timing(self, on):
        check(lib().ssb200_set_timing(self._h, C.c_int(1 if on else 0)))

    def timing(self):
        ms = (C.c_double * 2)()
        check(lib().ssb200_get_timing(self._h, ms))
        return dict(contraction_ms=ms[0], step_ms=ms[1])

    KERNEL_CLASSES = ("contraction", "step_all", "scan_stream", "scan_gram", "algebra", "step_mono")

    def kernel_profile(self):
        """Per kernel class of the lockstep loop: summed device ms (timing enabled) and signal-rows processed."""
        ms, rows, nbytes = (C.c_double * 6)(), (C.c_uint64 * 6)(), (C.c_uint64 * 8)()
        check(lib().ssb200_get_kernel_profile(self._h, ms, rows))
        check(lib().ssb200_get_kernel_bytes(self._h, nbytes))
        prof = {k: dict(ms=ms[i], rows=int(rows[i]), algorithmic_bytes=int(nbytes[i])) for i, k in enumerate(self.KERNEL_CLASSES)}
        prof["scan_gram"]["l2_bytes"] = int(nbytes[6])       # the K rows of G per signal
        return prof

    def transfer_bytes(self):
        """(host->device, device->host) bytes the host-pointer batch entry copied since the last reset_counters()."""
        b = (C.c_uint64 * 2)()
        check(lib().ssb200_get_transfer_bytes(self._h, b))
        return int(b[0]), int(b[1])

    def set_contraction_path(self, path):
        check(lib().ssb200_set_contraction_path(self._h, C.c_int({"auto": 0, "simt": 1, "tcgen05": 2, "tcgen05-tf32": 3, "fma": 4}[path])))


class Homotopy(_Solver):
    _name = "homotopy"

    def solve(self, b, tolerance=None, max_iterations=100):
        self._check_signal(b, 1)
        tol, mi = self._defaults(tolerance, max_iterations)
        x = np.empty(self.shape[1], dtype=self.dtype)
        it, err = C.c_uint32(), C.c_double()
        fn = getattr(lib(), f"ssb200_homotopy_solve_{self._sfx}")
        check(fn(self._h, _ptr(b), C.c_ssize_t(b.strides[0] // b.itemsize), tol, mi, _ptr(x), C.c_ssize_t(1),
                 C.byref(it), C.byref(err)))
        return x, HomotopyReport(it.value, err.value)

    def solve_batch(self, Y, tolerance=None, max_iterations=100, out=None, sparse=False):
        """Y: (B, M) rows are signals.  Returns X (B, N), iters (B,) uint32, solution_errors (B,) float64.

        sparse=True (host arrays too): the coefficients come back as (indices (B, cap) int32, values (B, cap), counts (B,) int32),
        cap = min(max_iterations + 1, N), entries beyond counts[b] unspecified -- no dense rows are formed anywhere.

        Y may also live on the device (torch CUDA tensor, or any DLPack / __cuda_array_interface__ producer): nothing is
        copied through the host, the work is queued on torch's current stream and X, iters (int32), solution_errors come
        back as torch tensors on that device.  With sparse=True the coefficients come back as (indices (B, cap) int32,
        values (B, cap), counts (B,) int32) with cap = min(max_iterations + 1, N) instead of the dense X."""
        if _is_cuda(Y):
            return self._solve_batch_cuda(_as_torch(Y), tolerance, max_iterations, out, sparse)
        self._check_signal(Y, 2)
        if Y.strides[1] != Y.itemsize:
            Y = np.ascontiguousarray(Y)
        tol, mi = self._defaults(tolerance, max_iterations)
        B = Y.shape[0]
        if sparse:
            if out is not None:
                raise ValueError("out= is for the dense result")
            cap = max(1, min(int(max_iterations) + 1, self.shape[1]))
            idx = np.empty((B, cap), dtype=np.int32)
            val = np.empty((B, cap), dtype=self.dtype)
            cnt = np.empty(B, dtype=np.int32)
            iters = np.empty(B, dtype=np.uint32)
            errs = np.empty(B, dtype=np.float64)
            check(getattr(lib(), f"ssb200_homotopy_solve_batch_sparse_{self._sfx}")(
                self._h, _ptr(Y), C.c_size_t(B), C.c_size_t(Y.strides[0] // Y.itemsize if B > 1 else self.shape[0]), tol, mi,
                _ptr(idx), _ptr(val), _ptr(cnt), C.c_size_t(cap), _ptr(iters), _ptr(errs)))
            return (idx, val, cnt), iters, errs
        X = out if out is not None else np.empty((B, self.shape[1]), dtype=self.dtype)
        if out is not None:      # the C side writes B rows of N values through this pointer: everything it assumes is checked here
            if not isinstance(X, np.ndarray) or X.dtype != self.dtype or X.shape != (B, self.shape[1]):
                raise ValueError(f"out must be a numpy array of dtype {self.dtype} and shape ({B}, {self.shape[1]})")
            if X.strides[1] != X.itemsize or (B > 1 and X.strides[0] < self.shape[1] * X.itemsize) or X.strides[0] % X.itemsize:
                raise ValueError("out must have unit inner stride and non-overlapping rows")
            if not X.flags.writeable:
                raise ValueError("out is read-only")
        iters = np.empty(B, dtype=np.uint32)
        errs = np.empty(B, dtype=np.float64)
        fn = getattr(lib(), f"ssb200_homotopy_solve_batch_{self._sfx}")
        check(fn(self._h, _ptr(Y), C.c_size_t(B), C.c_size_t(Y.strides[0] // Y.itemsize if B > 1 else self.shape[0]), tol, mi,
                 _ptr(X), C.c_size_t(X.strides[0] // X.itemsize if B > 1 else self.shape[1]), _ptr(iters), _ptr(errs)))
        return X, iters, errs

    def _solve_batch_cuda(self, Y, tolerance, max_iterations, out, sparse):
        import torch
        Y = self._check_cuda_signals(Y)
        B, N = Y.shape[0], self.shape[1]
        with torch.cuda.device(self.device):
            X = out if out is not None else torch.empty((B, N), dtype=Y.dtype, device=Y.device)
            if X.shape != (B, N) or X.dtype != Y.dtype or X.stride(1) != 1 or not X.is_cuda:
                raise ValueError("out must be a (batch, N) CUDA tensor of the signals' dtype with unit inner stride")
            iters = torch.empty(B, dtype=torch.int32, device=Y.device)
            errs = torch.empty(B, dtype=torch.float64, device=Y.device)
            stream = torch.cuda.current_stream().cuda_stream
            if B:
                self.solve_batch_ptr(Y.data_ptr(), B, Y.stride(0), X.data_ptr(), X.stride(0), iters.data_ptr(), errs.data_ptr(),
                                     tolerance, max_iterations, True, stream)
            if not sparse:
                return X, iters, errs
            cap = max(1, min(int(max_iterations) + 1, N))
            idx = torch.empty((B, cap), dtype=torch.int32, device=Y.device)
            val = torch.empty((B, cap), dtype=Y.dtype, device=Y.device)
            cnt = torch.empty(B, dtype=torch.int32, device=Y.device)
            check(getattr(lib(), f"ssb200_compress_rows_dev_{self._sfx}")(
                C.c_void_p(X.data_ptr()), C.c_size_t(B), C.c_size_t(N), C.c_size_t(X.stride(0)), C.c_size_t(cap),
                C.c_void_p(idx.data_ptr()), C.c_void_p(val.data_ptr()), C.c_void_p(cnt.data_ptr()), C.c_int(self.device),
                C.c_void_p(stream)))
            return (idx, val, cnt), iters, errs

    def solve_batch_ptr(self, y_ptr, batch, ldy, x_ptr, ldx, iters_ptr, errs_ptr, tolerance, max_iterations,
                        device_pointers, stream=0):
        """Raw-pointer entry (pinned host buffers or device buffers owned by the caller, e.g. torch tensors)."""
        tol, mi = self._defaults(tolerance, max_iterations)
        if device_pointers:
            fn = getattr(lib(), f"ssb200_homotopy_solve_batch_dev_{self._sfx}")
            check(fn(self._h, C.c_void_p(y_ptr), C.c_size_t(batch), C.c_size_t(ldy), tol, mi, C.c_void_p(x_ptr),
                     C.c_size_t(ldx), C.c_void_p(iters_ptr), C.c_void_p(errs_ptr), C.c_void_p(stream)))
        else:
            fn = getattr(lib(), f"ssb200_homotopy_solve_batch_{self._sfx}")
            check(fn(self._h, C.c_void_p(y_ptr), C.c_size_t(batch), C.c_size_t(ldy), tol, mi, C.c_void_p(x_ptr),
                     C.c_size_t(ldx), C.c_void_p(iters_ptr), C.c_void_p(errs_ptr)))

    RESIDUAL_MODES = {"auto": 0, "exact": 1, "incremental": 2}

    def set_residual_mode(self, mode):
        """How c = A^T (y - A x) is kept (include/ssb200.h): "exact" re-forms it every iteration as the reference does
        (homotopy-cpu.cpp:87-98, :255), "incremental" carries c -= gamma q, "auto" (default) picks by the dictionary's coherence."""
        check(lib().ssb200_homotopy_set_residual_mode(self._h, C.c_int(self.RESIDUAL_MODES[mode])))

    def residual_info(self):
        """(mode in effect: "exact" | "incremental", mutual coherence of the dictionary)"""
        mode, mu = C.c_int(), C.c_double()
        check(lib().ssb200_homotopy_get_residual_info(self._h, C.byref(mode), C.byref(mu)))
        return ("exact" if mode.value == 1 else "incremental"), mu.value

    def debug_contract(self, P, path="simt", reps=1):
        """Test/bench hook: Q[r] = A^T P[r] through one contraction kernel; returns (Q, ms per launch)."""
        P = np.ascontiguousarray(P, dtype=self.dtype)
        Q = np.empty((P.shape[0], self.shape[1]), dtype=self.dtype)
        ms = C.c_double()
        check(getattr(lib(), f"ssb200_debug_contract_{self._sfx}")(
            self._h, _ptr(P), C.c_size_t(P.shape[0]), _ptr(Q), C.c_int({"simt": 1, "tcgen05": 2, "tcgen05-tf32": 3, "fma": 4}[path]), C.c_int(reps),
            C.byref(ms)))
        return Q, ms.value

    def refresh(self, A):
        """Re-uploads an edited dictionary of the same shape (the reference state is a non-owning view,
        include/ss/policies.h:42; its bench edits A in place between solves, homotopy_bench.cpp:38-48)."""
        A = np.asarray(A)
        if A.shape != self.shape or A.dtype != self.dtype:
            raise ValueError("refresh must keep the dictionary's shape and dtype")
        rs, cs = A.strides[0] // A.itemsize, A.strides[1] // A.itemsize
        check(getattr(lib(), f"ssb200_homotopy_refresh_{self._sfx}")(self._h, _ptr(A), C.c_ssize_t(rs), C.c_ssize_t(cs)))

    def reconstruct(self, X):
        """Y[b] = A X[b] on the device (ss::reconstruct_signal, include/ss/ss.h:79-86)."""
        if _is_cuda(X):
            import torch
            X = _as_torch(X)
            X = X.reshape(1, -1) if X.ndim == 1 else X
            want = torch.float32 if self.dtype == np.float32 else torch.float64
            if X.dtype != want or X.shape[1] != self.shape[1] or X.device.index != self.device:
                raise ValueError(f"coefficients must be a (batch, {self.shape[1]}) {want} tensor on cuda:{self.device}")
            X = X if X.stride(1) == 1 else X.contiguous()
            with torch.cuda.device(self.device):
                Y = torch.empty((X.shape[0], self.shape[0]), dtype=X.dtype, device=X.device)
                check(getattr(lib(), f"ssb200_reconstruct_batch_dev_{self._sfx}")(
                    self._h, C.c_void_p(X.data_ptr()), C.c_size_t(X.shape[0]), C.c_size_t(X.stride(0)), C.c_void_p(Y.data_ptr()),
                    C.c_size_t(Y.stride(0)), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            return Y
        X = np.ascontiguousarray(np.atleast_2d(X), dtype=self.dtype)
        Y = np.empty((X.shape[0], self.shape[0]), dtype=self.dtype)
        check(getattr(lib(), f"ssb200_reconstruct_batch_{self._sfx}")(
            self._h, _ptr(X), C.c_size_t(X.shape[0]), C.c_size_t(X.shape[1]), _ptr(Y), C.c_size_t(Y.shape[1])))
        return Y


class HomotopyGroup:
    """One process, several GPUs of the box (ssb200_group_*, include/ssb200.h): the dictionary is replicated with one NCCL
    broadcast, solve_batch cuts the batch into contiguous shards, one host thread per GPU; results come back in the caller's
    arrays.  New functionality (the reference solves one signal per call on the CPU, include/ss/ss.h:50)."""

    @staticmethod
    def _preload_nccl():
        """The library binds NCCL at run time by soname.  torch ships its own (newer) libnccl.so.2: if the system copy were loaded
        first, a later `import torch` would resolve to it and fail on missing symbols -- so torch's copy, when installed, goes first."""
        import importlib.util
        import os
        try:
            spec = importlib.util.find_spec("nvidia.nccl")
            for d in (spec.submodule_search_locations if spec else []):
                p = os.path.join(d, "lib", "libnccl.so.2")
                if os.path.exists(p):
                    C.CDLL(p, mode=C.RTLD_GLOBAL)
                    return
        except (ImportError, OSError, ValueError):
            pass

    def __init__(self, A, devices=None):
        A = np.asarray(A)
        if A.ndim != 2:
            raise ValueError("A must be 2-dimensional")
        if A.dtype not in _SFX:
            raise TypeError("A must be float32 or float64")
        self._preload_nccl()
        if devices is None:
            devices = list(range(max(1, lib().ssb200_device_count())))
        self.devices = [int(d) for d in devices]
        self._sfx, self._ct = _SFX[A.dtype]
        self.dtype, self.shape = A.dtype, A.shape
        self._g = C.c_void_p()
        devs = (C.c_int * len(self.devices))(*self.devices)
        rs, cs = A.strides[0] // A.itemsize, A.strides[1] // A.itemsize
        check(getattr(lib(), f"ssb200_group_create_{self._sfx}")(_ptr(A), C.c_size_t(A.shape[0]), C.c_size_t(A.shape[1]), C.c_ssize_t(rs),
                                                                 C.c_ssize_t(cs), devs, C.c_int(len(self.devices)), C.byref(self._g)))

    def __del__(self):
        g = getattr(self, "_g", None)
        if g is not None and g.value:
            lib().ssb200_group_destroy(g)
            self._g = C.c_void_p()

    def __len__(self):
        return int(lib().ssb200_group_size(self._g))

    def set_residual_mode(self, mode):
        check(lib().ssb200_group_set_residual_mode(self._g, C.c_int(Homotopy.RESIDUAL_MODES[mode])))

    def solve_batch(self, Y, tolerance=None, max_iterations=100):
        """Y: (B, M) host array or a CUDA tensor on the group's first device.  Returns X, iters, solution_errors of the same kind."""
        tol = self._ct(10 * np.finfo(self.dtype).eps if tolerance is None else tolerance)
        mi = C.c_uint32(int(max_iterations))
        if _is_cuda(Y):
            import torch
            Y = _as_torch(Y)
            want = torch.float32 if self.dtype == np.float32 else torch.float64
            if Y.dtype != want or Y.ndim != 2 or Y.shape[1] != self.shape[0] or Y.device.index != self.devices[0]:
                raise ValueError(f"signals must be a (batch, {self.shape[0]}) {want} tensor on cuda:{self.devices[0]}")
            Y = Y.contiguous()
            B = Y.shape[0]
            with torch.cuda.device(self.devices[0]):
                X = torch.empty((B, self.shape[1]), dtype=Y.dtype, device=Y.device)
                iters = torch.empty(B, dtype=torch.int32, device=Y.device)
                errs = torch.empty(B, dtype=torch.float64, device=Y.device)
                check(getattr(lib(), f"ssb200_group_solve_batch_dev_{self._sfx}")(
                    self._g, C.c_void_p(Y.data_ptr()), C.c_size_t(B), C.c_size_t(self.shape[0]), tol, mi, C.c_void_p(X.data_ptr()),
                    C.c_size_t(self.shape[1]), C.c_void_p(iters.data_ptr()), C.c_void_p(errs.data_ptr()),
                    C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            return X, iters, errs
        if not isinstance(Y, np.ndarray) or Y.dtype != self.dtype:
            raise TypeError(f"signals must be a numpy array of dtype {self.dtype}")
        if Y.ndim != 2 or Y.shape[1] != self.shape[0]:
            raise ValueError(f"signals must be (batch, {self.shape[0]})")
        Y = np.ascontiguousarray(Y)
        B = Y.shape[0]
        X = np.empty((B, self.shape[1]), dtype=self.dtype)
        iters = np.empty(B, dtype=np.uint32)
        errs = np.empty(B, dtype=np.float64)
        check(getattr(lib(), f"ssb200_group_solve_batch_{self._sfx}")(
            self._g, _ptr(Y), C.c_size_t(B), C.c_size_t(self.shape[0]), tol, mi, _ptr(X), C.c_size_t(self.shape[1]), _ptr(iters), _ptr(errs)))
        return X, iters, errs


class Irls(_Solver):
    _name = "irls"

    def solve(self, b, tolerance=None, max_iterations=100):
        self._check_signal(b, 1)
        tol, mi = self._defaults(tolerance, max_iterations)
        x = np.empty(self.shape[1], dtype=self.dtype)
        it, err, spd = C.c_uint32(), C.c_double(), C.c_uint8()
        fn = getattr(lib(), f"ssb200_irls_solve_{self._sfx}")
        check(fn(self._h, _ptr(b), C.c_ssize_t(b.strides[0] // b.itemsize), tol, mi, _ptr(x), C.c_ssize_t(1),
                 C.byref(it), C.byref(err), C.byref(spd)))
        return x, IrlsReport(it.value, err.value, bool(spd.value))

    def solve_batch(self, Y, tolerance=None, max_iterations=100):
        self._check_signal(Y, 2)
        if Y.strides[1] != Y.itemsize:
            Y = np.ascontiguousarray(Y)
        tol, mi = self._defaults(tolerance, max_iterations)
        B = Y.shape[0]
        X = np.empty((B, self.shape[1]), dtype=self.dtype)
        iters = np.empty(B, dtype=np.uint32)
        errs = np.empty(B, dtype=np.float64)
        spd = np.empty(B, dtype=np.uint8)
        fn = getattr(lib(), f"ssb200_irls_solve_batch_{self._sfx}")
        check(fn(self._h, _ptr(Y), C.c_size_t(B), C.c_size_t(Y.strides[0] // Y.itemsize), tol, mi,
                 _ptr(X), C.c_size_t(X.shape[1]), _ptr(iters), _ptr(errs), _ptr(spd)))
        return X, iters, errs, spd.astype(bool)


def fp64_peak(kind="dmma", device=0):
    """Measured FP64 peak of the device in TFLOP/s: independent DFMA chains ("dfma") or mma.sync.m8n8k4.f64 tiles ("dmma")."""
    t = C.c_double()
    check(lib().ssb200_fp64_peak_probe(C.c_int(device), C.c_int({"dfma": 0, "dmma": 1}[kind]), C.byref(t)))
    return t.value


def reconstruct_signal(A, x, device=0):
    """ss::reconstruct_signal (include/ss/ss.h:79-86): y = A x."""
    return Homotopy(A, device).reconstruct(x)[0]


def norm_l1(A, device=0):
    """ss::norm_l1 (include/ss/ss.h:88-93): divides every column by its l1 norm, in place."""
    if A.dtype not in _SFX or A.ndim != 2:
        raise TypeError("A must be a 2-D float32/float64 array")
    sfx, _ = _SFX[A.dtype]
    rs, cs = A.strides[0] // A.itemsize, A.strides[1] // A.itemsize
    check(getattr(lib(), f"ssb200_norm_l1_{sfx}")(_ptr(A), C.c_size_t(A.shape[0]), C.c_size_t(A.shape[1]),
                                                   C.c_ssize_t(rs), C.c_ssize_t(cs), C.c_int(device)))
    return A


NOALLOC_SCRATCH_BYTES = 1 << 20   # include/ssb200.h: SSB200_NOALLOC_SCRATCH_BYTES


def norm_l1_noalloc(dA, stream=None):
    """ss::norm_l1 in place on a DEVICE matrix (torch CUDA tensor, 2-D, unit column stride) through the entry that allocates
    nothing (ssb200_norm_l1_dev_noalloc_*; no counterpart in the reference).  The column sums live in a static device array of
    NOALLOC_SCRATCH_BYTES, so at most that many bytes of columns; queued on `stream` (default: torch's current stream) and NOT
    synchronised; calls that overlap on one GPU race on the scratch array."""
    import torch
    dt = {torch.float32: np.dtype(np.float32), torch.float64: np.dtype(np.float64)}.get(dA.dtype)
    if dt is None or dA.dim() != 2 or not dA.is_cuda or dA.stride(1) != 1:
        raise TypeError("dA must be a 2-D float32/float64 CUDA tensor with unit column stride")
    sfx, _ = _SFX[dt]
    if stream is None:
        stream = torch.cuda.current_stream(dA.device).cuda_stream
    check(getattr(lib(), f"ssb200_norm_l1_dev_noalloc_{sfx}")(C.c_void_p(dA.data_ptr()), C.c_size_t(dA.shape[0]), C.c_size_t(dA.shape[1]),
                                                               C.c_size_t(dA.stride(0)), C.c_int(dA.device.index or 0), C.c_void_p(stream)))
    return dA
