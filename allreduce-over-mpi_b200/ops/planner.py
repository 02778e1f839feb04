"""Python face of the native planner (csrc/planner). No GPU needed.

Capability parity (reference paths): cost_model/GetWidth.h (enumerate_widths), ChooseWidth.h
(candidates), CostModel.h (legacy_cost / legacy_choose), GetPrimeFactor.h / IsPrimeNumber.h,
PrintTreeStructure.h (format_widths), topo_count/factor_count.py (count_factorizations),
allreduce_over_mpi/mpi_mod.hpp:882-929 (parse_topo), :107-144 (print_ops).
"""
import ctypes as C

from .. import _lib

__all__ = ["count_factorizations", "enumerate_widths", "candidates", "prime_factors", "is_prime",
           "is_prime_legacy", "legacy_cost", "legacy_choose", "plan_traffic", "print_ops", "LinkModel",
           "parse_topo", "parse_topo_ex", "format_widths", "version"]


def _arr(ws):
    ws = list(ws)
    return (C.c_int * max(1, len(ws)))(*ws), len(ws)


def version():
    return _lib.load().ft_version_string().decode()


def count_factorizations(n):
    """H(n): number of ordered factorizations (H(0)=0, H(1)=1)."""
    return int(_lib.load().ft_count_factorizations(int(n)))


def enumerate_widths(n, legacy_flat=False):
    L = _lib.load()
    nl = C.c_longlong(0)
    need = -L.ft_enumerate_widths(n, int(legacy_flat), None, 0, C.byref(nl))
    if need <= 0:
        return [[]] if nl.value == 1 else []
    buf = (C.c_int * need)()
    L.ft_enumerate_widths(n, int(legacy_flat), buf, need, C.byref(nl))
    out, i = [], 0
    while i < need:
        k = buf[i]
        out.append(list(buf[i + 1:i + 1 + k]))
        i += 1 + k
    return out


def candidates(n):
    """Reference ``chooseWidth(n)`` listing as strings, e.g. ``['2*4', '3*3-1', '7*1+1']``."""
    L = _lib.load()
    need = L.ft_candidates_text(n, None, 0)
    buf = C.create_string_buffer(need + 1)
    L.ft_candidates_text(n, buf, need + 1)
    return [s for s in buf.value.decode().split("\n") if s]


def prime_factors(n):
    L = _lib.load()
    buf = (C.c_int * 64)()
    k = L.ft_prime_factors(n, buf, 64)
    return list(buf[:k])


def is_prime(n):
    return bool(_lib.load().ft_is_prime(n))


def is_prime_legacy(n):
    return bool(_lib.load().ft_is_prime_legacy(n))


def format_widths(widths, lonely=0):
    s = "*".join(str(w) for w in widths)
    if lonely > 0:
        s += f"+{lonely}"
    elif lonely < 0:
        s += f"{lonely}"
    return s


def legacy_cost(widths, total_nodes, chunk=100.0, terms=False):
    a, k = _arr(widths)
    t = (C.c_double * 3)()
    c = _lib.load().ft_legacy_cost(a, k, total_nodes, float(chunk), t)
    return (c, tuple(t)) if terms else c


def legacy_choose(n, chunk=100.0):
    buf = (C.c_int * 64)()
    cost = C.c_double(0)
    k = _lib.load().ft_legacy_choose(n, float(chunk), buf, 64, C.byref(cost))
    return list(buf[:k]), cost.value


def plan_traffic(widths, rank, count):
    a, k = _arr(widths)
    m, e = C.c_longlong(0), C.c_longlong(0)
    rc = _lib.load().ft_plan_traffic(a, k, rank, count, C.byref(m), C.byref(e))
    if rc:
        raise ValueError(f"bad plan {widths}")
    return m.value, e.value


def print_ops(widths, rank, lonely=0):
    L = _lib.load()
    a, k = _arr(widths)
    need = L.ft_print_ops(a, k, rank, lonely, None, 0)
    if need < 0:
        raise ValueError(f"bad plan {widths} / rank {rank}")
    buf = C.create_string_buffer(need + 1)
    L.ft_print_ops(a, k, rank, lonely, buf, need + 1)
    return buf.value.decode()


def parse_topo(text, nranks):
    """FT_TOPO semantics of the reference: '' -> [n]; any 1 -> [1] (ring); bad product -> ValueError."""
    buf = (C.c_int * 64)()
    k = _lib.load().ft_parse_topo(None if text is None else text.encode(), nranks, buf, 64)
    if k < 0:
        raise ValueError(_lib.load().ft_last_error().decode())
    return list(buf[:k])


def parse_topo_ex(text, nranks):
    """FT_TOPO with the tree-size suffixes: '2,3+1' (one lonely rank), '2,4-1' (one phantom position).
    Returns (widths, lonely, phantoms)."""
    buf = (C.c_int * 64)()
    lonely, phantoms = C.c_int(0), C.c_int(0)
    k = _lib.load().ft_parse_topo_ex(None if text is None else text.encode(), nranks, buf, 64, C.byref(lonely), C.byref(phantoms))
    if k < 0:
        raise ValueError(_lib.load().ft_last_error().decode())
    return list(buf[:k]), lonely.value, phantoms.value


class LinkModel:
    """alpha-beta-gamma model of the box; predicts every algorithm, picks the best one."""

    def __init__(self, json_path=None):
        self._L = _lib.load()
        self._h = self._L.ft_link_model_create(json_path.encode() if json_path else None)
        if not self._h:
            raise ValueError(self._L.ft_last_error().decode())

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.ft_link_model_destroy(self._h)
            self._h = None

    def set(self, **kw):
        for k, v in kw.items():
            if self._L.ft_link_model_set(self._h, k.encode(), float(v)):
                raise KeyError(k)
        return self

    def dump(self):
        import json
        buf = C.create_string_buffer(2048)
        self._L.ft_link_model_dump(self._h, buf, 2048)
        return json.loads(buf.value.decode())

    def predict_us(self, algo, widths, nranks, nbytes, registered=False):
        a, k = _arr(widths)
        return self._L.ft_predict_us(self._h, int(algo), a, k, nranks, float(nbytes), int(registered))

    def choose(self, nranks, nbytes, registered=False):
        buf = (C.c_int * 16)()
        algo, est = C.c_int(0), C.c_double(0)
        k = self._L.ft_choose(self._h, nranks, float(nbytes), int(registered), C.byref(algo), buf, 16, C.byref(est))
        return algo.value, list(buf[:k]), est.value
