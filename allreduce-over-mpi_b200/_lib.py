"""ctypes loader for the native libraries.

``libflextree.so`` (nvcc build: planner + CPU world + CUDA runtime/kernels) is preferred; on a machine
where it has not been built or cannot be loaded the g++-only ``libflextree_host.so`` (planner + CPU
world) is used so that the planner and the CPU data plane keep working. On a machine WITH a GPU the
CUDA library is mandatory: we fail loudly instead of silently falling back.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(_HERE)
CUDA_LIB = os.path.join(_HERE, "lib", "libflextree.so")
HOST_LIB = os.path.join(REPO, "build", "libflextree_host.so")

_lib = None
_has_cuda_lib = False


def _gpu_present():
    return os.path.exists("/dev/nvidiactl") or os.path.exists("/dev/nvidia0")


def build(target="all", jobs=None):
    """Run the in-tree Makefile (``host``, ``cuda``, ``tools`` or ``all``)."""
    jobs = jobs or os.cpu_count() or 4
    subprocess.check_call(["make", "-C", REPO, f"-j{jobs}", target])


def load():
    global _lib, _has_cuda_lib
    if _lib is not None:
        return _lib
    err = None
    if os.path.exists(CUDA_LIB):
        try:
            _lib = C.CDLL(CUDA_LIB, mode=C.RTLD_GLOBAL)
            _has_cuda_lib = True
        except OSError as e:  # e.g. built for another image
            err = e
    if _lib is None:
        if _gpu_present() and os.environ.get("FT_ALLOW_HOST_ONLY") != "1":
            raise RuntimeError(
                f"flextree: GPU present but {CUDA_LIB} is missing or unloadable ({err}); "
                "run `make cuda` (or `python -c 'import __graft_entry__ as g; g.build()'`)")
        if not os.path.exists(HOST_LIB):
            build("host")
        _lib = C.CDLL(HOST_LIB, mode=C.RTLD_GLOBAL)
    _declare(_lib)
    return _lib


_fast = None
_fast_tried = False


def fast():
    """The compiled per-call module (csrc/python/ftfast.c -> lib/_ftfast.so), or None when it is not built."""
    global _fast, _fast_tried
    if _fast_tried:
        return _fast
    _fast_tried = True
    load()
    path = os.path.join(_HERE, "lib", "_ftfast.so")
    if _has_cuda_lib and os.path.exists(path) and os.environ.get("FT_NO_FASTCALL") != "1":
        try:
            import importlib.machinery
            import importlib.util
            loader = importlib.machinery.ExtensionFileLoader("_ftfast", path)
            spec = importlib.util.spec_from_loader("_ftfast", loader)
            mod = importlib.util.module_from_spec(spec)
            loader.exec_module(mod)
            _fast = mod
        except Exception:  # e.g. built for another Python: the ctypes path still works
            _fast = None
    return _fast


def has_cuda_lib():
    load()
    return _has_cuda_lib


def _declare(L):
    i32, i64, u64, dbl, vp, cp = C.c_int, C.c_longlong, C.c_uint64, C.c_double, C.c_void_p, C.c_char_p
    pi32 = C.POINTER(C.c_int)

    def sig(name, res, args):
        f = getattr(L, name, None)
        if f is not None:
            f.restype = res
            f.argtypes = args

    sig("ft_last_error", cp, [])
    sig("ft_version_string", cp, [])
    sig("ft_count_factorizations", u64, [u64])
    sig("ft_enumerate_widths", i64, [i32, i32, pi32, i64, C.POINTER(i64)])
    sig("ft_candidates_text", i64, [i32, cp, i64])
    sig("ft_prime_factors", i32, [i32, pi32, i32])
    sig("ft_is_prime", i32, [i32])
    sig("ft_is_prime_legacy", i32, [i32])
    sig("ft_legacy_cost", dbl, [pi32, i32, i32, dbl, C.POINTER(dbl)])
    sig("ft_legacy_choose", i32, [i32, dbl, pi32, i32, C.POINTER(dbl)])
    sig("ft_plan_traffic", i32, [pi32, i32, i32, i64, C.POINTER(i64), C.POINTER(i64)])
    sig("ft_print_ops", i64, [pi32, i32, i32, i32, cp, i64])
    sig("ft_link_model_create", vp, [cp])
    sig("ft_link_model_destroy", None, [vp])
    sig("ft_link_model_dump", i64, [vp, cp, i64])
    sig("ft_link_model_set", i32, [vp, cp, dbl])
    sig("ft_predict_us", dbl, [vp, i32, pi32, i32, i32, dbl, i32])
    sig("ft_choose", i32, [vp, i32, dbl, i32, pi32, pi32, i32, C.POINTER(dbl)])
    sig("ft_parse_topo", i32, [cp, i32, pi32, i32])
    sig("ft_parse_topo_ex", i32, [cp, i32, pi32, i32, pi32, pi32])
    # CPU world
    sig("ft_host_world_create", vp, [i32, C.c_size_t, C.c_size_t])
    sig("ft_host_world_destroy", None, [vp])
    sig("ft_host_registered", vp, [vp, i32])
    sig("ft_host_error", C.c_uint, [vp, i32])
    sig("ft_host_clear_error", None, [vp])
    sig("ft_host_epoch", C.c_ulonglong, [vp, i32])
    sig("ft_host_allreduce", i64, [vp, C.POINTER(vp), C.POINTER(vp), i64, i32, i32, i32, i32, pi32, i32,
                                   i32, i64, i32, dbl, i32, C.c_ulonglong, i32, i32, i32])
    sig("ft_bf16_to_float", C.c_float, [C.c_uint16])
    sig("ft_float_to_bf16", C.c_uint16, [C.c_float])
    sig("ft_f16_to_float", C.c_float, [C.c_uint16])
    sig("ft_float_to_f16", C.c_uint16, [C.c_float])
    sig("ft_e4m3_to_float", C.c_float, [C.c_uint])
    sig("ft_float_to_e4m3", C.c_uint, [C.c_float])
    sig("ft_fp8_block_scale", C.c_uint, [C.c_float])
    # CUDA runtime (absent from the host-only library)
    sig("ft_comm_options_default", None, [vp])
    sig("ft_comm_create", i32, [C.POINTER(vp), i32, i32, i32, vp, vp, vp])
    sig("ft_comm_create_local", i32, [C.POINTER(vp), i32, pi32, vp])
    sig("ft_comm_create_emulated", i32, [C.POINTER(vp), i32, i32, vp])
    sig("ft_comm_destroy", i32, [vp])
    sig("ft_comm_rank", i32, [vp])
    sig("ft_comm_size", i32, [vp])
    sig("ft_comm_has_multicast", i32, [vp])
    sig("ft_comm_error", C.c_uint, [vp])
    sig("ft_comm_clear_error", i32, [vp])
    sig("ft_comm_error_peek", C.c_uint, [vp])
    sig("ft_comm_inject_fault", i32, [vp, i32])
    sig("ft_comm_set_tuning", i32, [vp, cp, i64])
    sig("ft_link_probe", i32, [vp, i32, i32, C.c_size_t, i32, vp])
    sig("ft_comm_set_widths", i32, [vp, pi32, i32])
    sig("ft_comm_set_algo", i32, [vp, i32])
    sig("ft_comm_set_link_model", i32, [vp, vp])
    sig("ft_comm_set_launch", i32, [vp, i32, i32, i64])
    sig("ft_comm_set_tma", i32, [vp, i32])
    sig("ft_comm_set_max_ctas", i32, [vp, i32])
    sig("ft_comm_trace_enable", i32, [vp, i32])
    sig("ft_comm_trace_read", i32, [vp, i32, C.POINTER(C.c_ulonglong), i32])
    sig("ft_comm_reserve", i32, [vp, C.c_size_t])
    sig("ft_symm_alloc", i32, [vp, C.c_size_t, C.POINTER(vp)])
    sig("ft_symm_free", i32, [vp, vp])
    sig("ft_allreduce", i32, [vp, vp, C.c_size_t, i32, i32, vp, vp])
    sig("ft_allreduce_ex", i32, [vp, vp, C.c_size_t, i32, i32, i32, dbl, vp, vp])
    sig("ft_allreduce_multi", i32, [C.POINTER(vp), C.POINTER(vp), C.c_size_t, i32, i32, i32, dbl, vp, vp])
    sig("ft_comm_last_call", i32, [vp, vp])
    sig("ft_allreduce_host", i32, [vp, vp, C.c_size_t, i32, i32, vp])
    sig("ft_allreduce_host_multi", i32, [C.POINTER(vp), C.POINTER(vp), C.c_size_t, i32, i32, dbl, vp, C.c_size_t])
    sig("ft_reduce_scatter", i32, [vp, vp, C.c_size_t, i32, i32, vp, vp])
    sig("ft_all_gather", i32, [vp, vp, C.c_size_t, i32, vp, vp])
    sig("ft_reduce_scatter_multi", i32, [C.POINTER(vp), C.POINTER(vp), C.c_size_t, i32, i32, vp, vp])
    sig("ft_all_gather_multi", i32, [C.POINTER(vp), C.POINTER(vp), C.c_size_t, i32, vp, vp])
    sig("ft_allreduce_emulated_noalloc", i32, [C.POINTER(vp), C.POINTER(vp), i32, C.c_size_t, i32, i32, vp])
    sig("ft_noalloc_error", C.c_uint, [i32])
