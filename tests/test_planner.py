"""Planner unit tests against values captured from the unmodified reference."""
import pytest

import flextree as ft

H_0_32 = [0, 1, 1, 1, 2, 1, 3, 1, 4, 2, 3, 1, 8, 1, 3, 3, 8, 1, 8, 1, 8, 3, 3, 1, 20, 2, 3, 4, 8, 1, 13, 1, 16]


def fmt(lists):
    return [ft.format_widths(w) for w in lists]


def test_count_factorizations_oracle():
    assert [ft.count_factorizations(n) for n in range(33)] == H_0_32
    assert ft.count_factorizations(64) == 32
    assert ft.count_factorizations(128) == 64
    assert ft.count_factorizations(41160) == 24924


def test_count_matches_python_recurrence():
    def h(n):
        if n == 0:
            return 0
        if n == 1:
            return 1
        return sum(h(n // d) for d in range(2, n + 1) if n % d == 0)
    for n in range(0, 200):
        assert ft.count_factorizations(n) == h(n)


@pytest.mark.parametrize("n,expect", [
    (2, ["1*2", "2*1"]),
    (4, ["2*2", "1*4", "4*1"]),
    (6, ["2*3", "3*2", "1*6", "6*1"]),
    (8, ["2*2*2", "2*4", "4*2", "1*8", "8*1"]),
    (12, ["2*2*3", "2*3*2", "2*6", "3*2*2", "3*4", "4*3", "6*2", "1*12", "12*1"]),
])
def test_enumerate_legacy_listing(n, expect):
    assert fmt(ft.enumerate_widths(n, legacy_flat=True)) == expect


def test_enumerate_edge_cases():
    assert ft.enumerate_widths(1) == [[]]
    assert ft.enumerate_widths(0) == []
    assert ft.enumerate_widths(7) == [[7]]


@pytest.mark.parametrize("n,cnt", [(2, 2), (3, 2), (4, 3), (5, 2), (6, 4), (7, 2), (8, 5), (12, 9),
                                   (16, 9), (24, 21), (32, 17), (64, 33)])
def test_enumerate_counts(n, cnt):
    legacy = ft.enumerate_widths(n, legacy_flat=True)
    assert len(legacy) == cnt == ft.count_factorizations(n) + 1
    plain = ft.enumerate_widths(n)
    assert len(plain) == ft.count_factorizations(n)
    for w in plain:
        p = 1
        for x in w:
            assert x >= 2
            p *= x
        assert p == n
    assert len({tuple(w) for w in plain}) == len(plain)


def test_candidates_oracle():
    assert ft.candidates(7) == ["2*2*2-1", "2*4-1", "4*2-1", "1*8-1", "8*1-1", "2*3+1", "3*2+1", "1*6+1", "6*1+1"]
    assert ft.candidates(8) == ["2*2*2", "2*4", "4*2", "1*8", "8*1", "3*3-1", "1*9-1", "9*1-1", "1*7+1", "7*1+1"]


def test_primes():
    assert ft.prime_factors(41160) == [2, 2, 2, 3, 5, 7, 7, 7]
    assert ft.prime_factors(1) == []
    assert [n for n in range(30) if ft.is_prime(n)] == [2, 3, 5, 7, 11, 13, 17, 19, 23, 29]
    assert ft.is_prime_legacy(1) and not ft.is_prime(1)  # reference quirk kept only in the legacy variant
    assert not ft.is_prime_legacy(0)


def test_legacy_cost_oracle():
    six = {"2*3": 0.656, "3*2": 0.669333, "1*6": 0.642667, "6*1": 0.709333}
    for w in ft.enumerate_widths(6, True):
        assert ft.legacy_cost(w, 6, 100) == pytest.approx(six[ft.format_widths(w)], abs=1e-6)
    eight = {"2*2*2": 0.724, "2*4": 0.676, "4*2": 0.696, "1*8": 0.666, "8*1": 0.736}
    for w in ft.enumerate_widths(8, True):
        assert ft.legacy_cost(w, 8, 100) == pytest.approx(eight[ft.format_widths(w)], abs=1e-6)
    c, (lat, mem, bw) = ft.legacy_cost([2, 4], 8, 100, terms=True)
    assert lat == pytest.approx(0.016) and mem == pytest.approx(0.065) and bw == pytest.approx(0.595)


@pytest.mark.parametrize("chunk,table", [
    (100, {6: "1*6", 8: "1*8", 9: "1*9", 10: "2*5", 12: "2*6", 16: "2*8", 20: "4*5", 24: "3*8", 32: "4*8",
           64: "8*8", 128: "2*8*8", 256: "4*8*8"}),
    (1, {8: "1*8", 10: "2*5", 64: "8*8"}),
    (10000, {8: "1*8", 24: "1*24", 32: "2*16", 64: "4*16", 128: "8*16", 256: "8*32"}),
])
def test_legacy_choose_oracle(chunk, table):
    for n, want in table.items():
        w, _ = ft.legacy_choose(n, chunk)
        assert ft.format_widths(w) == want, (n, chunk)


def test_legacy_cost_any_height():
    # the reference has no return for heights 0 or > 9; ours is total
    assert ft.legacy_cost([2] * 10, 1024, 100) > 0
    assert ft.legacy_cost([], 1, 100) >= 0


def test_parse_topo():
    assert ft.parse_topo("2,3", 6) == [2, 3]
    assert ft.parse_topo("2 3", 6) == [2, 3]
    assert ft.parse_topo("", 6) == [6]
    assert ft.parse_topo("   ", 8) == [8]
    assert ft.parse_topo("1", 6) == [1]          # ring
    assert ft.parse_topo("2,1,3", 6) == [1]      # any 1 -> ring
    with pytest.raises(ValueError):
        ft.parse_topo("2,2", 6)
    with pytest.raises(ValueError):
        ft.parse_topo("2,a", 6)


def test_parse_topo_env(monkeypatch):
    monkeypatch.setenv("FT_TOPO", "4,2")
    assert ft.parse_topo(None, 8) == [4, 2]
    monkeypatch.delenv("FT_TOPO")
    assert ft.parse_topo(None, 8) == [8]


def test_print_ops_oracle():
    txt = ft.print_ops([2, 3], 4)
    send, recv = txt.split("recv ops")
    assert "peer 4 (self): blocks 0 2 4" in send and "peer 5: blocks 1 3 5" in send
    assert "peer 0: blocks 0" in send and "peer 2: blocks 2" in send
    assert recv.count("blocks 0 2 4") == 2 and recv.count(": blocks 4\n") == 3
    txt = ft.print_ops([2, 2, 2], 5)
    send = txt.split("recv ops")[0]
    assert "peer 5 (self): blocks 1 5" in send and "peer 7: blocks 3 7" in send
    assert "peer 1: blocks 1" in send
    txt = ft.print_ops([4, 2], 5)
    send = txt.split("recv ops")[0]
    for line in ("peer 4: blocks 0 4", "peer 5 (self): blocks 1 5", "peer 6: blocks 2 6", "peer 7: blocks 3 7"):
        assert line in send


@pytest.mark.parametrize("n,widths_list,elems,msgs", [
    (6, [[2, 3], [3, 2], [6]], 200_000_000 // 4 // 5, 300 // 5),
    (8, [[8], [2, 4], [4, 2], [2, 2, 2]], 280_000_000 // 4 // 5, 560 // 5),
])
def test_traffic_is_factorization_invariant(n, widths_list, elems, msgs):
    """Every factorization moves the same bytes in the same number of block messages
    (reference: 200,000,000 B / 300 msgs for 5 x 1M floats at N=6; 280,000,000 B / 560 msgs at N=8)."""
    count = 1_000_000
    for w in widths_list:
        tot_m = tot_e = 0
        for r in range(n):
            m, e = ft.plan_traffic(w, r, count)
            tot_m += m
            tot_e += e
        assert tot_m == msgs and tot_e == elems, w


def test_traffic_small_count():
    # N=6, 35 floats: 60 messages / 1400 bytes in total (reference benchmark default shape)
    tot_m = tot_e = 0
    for r in range(6):
        m, e = ft.plan_traffic([2, 3], r, 35)
        tot_m += m
        tot_e += e
    assert tot_m == 60 and tot_e * 4 == 1400


def test_link_model_choice_is_sane():
    m = ft.LinkModel()
    algo, widths, est = m.choose(8, 1024)
    assert algo == ft.ALGO_ONESHOT
    algo, widths, est = m.choose(8, 256 << 20)
    assert algo == ft.ALGO_STAGES and eval("*".join(map(str, widths))) == 8
    algo, widths, est = m.choose(2, 256 << 20)
    assert algo == ft.ALGO_STAGES and widths == [2]
    m.set(has_nvls=1)
    algo, widths, est = m.choose(8, 256 << 20)
    assert algo == ft.ALGO_NVLS
    # the switch moves (1 + 1/N) S per link direction, the P2P stages 2 (N-1)/N S: no NVLS win at N = 2
    assert m.choose(2, 256 << 20)[0] == ft.ALGO_STAGES
    # more stages = more flag hops: deep trees lose at small sizes on a uniform fabric
    assert m.predict_us(ft.ALGO_STAGES, [2, 2, 2], 8, 1 << 16) > m.predict_us(ft.ALGO_STAGES, [8], 8, 1 << 16)
    # fitted constants reproduce the 8xB200 measurements of round 1 (profiles/cost_model_validation.md)
    for nbytes, us in ((4 << 20, 31.9), (16 << 20, 57.8), (64 << 20, 161.3), (256 << 20, 584.0), (1 << 30, 2246.1)):
        assert abs(m.predict_us(ft.ALGO_NVLS, [8], 8, nbytes, True) - us) / us < 0.20
    for nbytes, us in ((1 << 10, 13.8), (64 << 10, 13.4)):
        assert abs(m.predict_us(ft.ALGO_ONESHOT, [8], 8, nbytes, True) - us) / us < 0.20
    # 7 ranks (prime) still get an answer
    algo, widths, _ = m.choose(7, 1 << 24)
    assert widths in ([7], [1])


def test_link_model_json_roundtrip(tmp_path):
    import json
    p = tmp_path / "m.json"
    p.write_text(json.dumps({"alpha_sync_us": 3.5, "beta_p2p_GBps": 650.0, "has_nvls": True, "note": "x"}))
    m = ft.LinkModel(str(p))
    d = m.dump()
    assert d["alpha_sync_us"] == 3.5 and d["beta_p2p_GBps"] == 650.0 and d["has_nvls"] is True
    with pytest.raises(ValueError):
        ft.LinkModel(str(tmp_path / "missing.json"))


def test_legacy_enumerator_getwidth2_sets():
    """The reference's older enumerator (getWidth2, cost_model/GetWidth.h:51-227) yields the same SET, sorted:
    `getWidth2(8)` = 1*8, 2*2*2, 2*4, 4*2, 8*1 and `getWidth2(12)` = sorted(getWidth(12))."""
    assert sorted(fmt(ft.enumerate_widths(8, True))) == sorted(["1*8", "2*2*2", "2*4", "4*2", "8*1"])
    assert sorted(map(tuple, ft.enumerate_widths(8, True))) == [(1, 8), (2, 2, 2), (2, 4), (4, 2), (8, 1)]
    assert len(ft.enumerate_widths(12, True)) == 9
    # no 9-stage limit here (getWidth2 caps the depth at 9): 2^12 has a 12-stage tree
    assert [2] * 12 in ft.enumerate_widths(4096)


def test_ft_wire_environment_default():
    """FT_WIRE: default wire format of the torch path when the caller passes none."""
    from flextree.ops import constants as K
    assert K.default_wire(K.FLOAT, "sum", {}) == K.FLOAT
    assert K.default_wire(K.FLOAT, "sum", {"FT_WIRE": "native"}) == K.FLOAT
    assert K.default_wire(K.FLOAT, "sum", {"FT_WIRE": "bf16"}) == K.BFLOAT16
    assert K.default_wire(K.BFLOAT16, "sum", {"FT_WIRE": "bf16"}) == K.BFLOAT16
    assert K.default_wire(K.INT32, "sum", {"FT_WIRE": "bf16"}) == K.INT32          # integers are never narrowed
    assert K.default_wire(K.FLOAT, "max", {"FT_WIRE": "bf16"}) == K.FLOAT          # only SUM
    assert K.default_wire(K.FLOAT, "sum", {"FT_WIRE": "fp8"}) == K.FP8_BLOCK
    assert K.default_wire(K.BFLOAT16, "sum", {"FT_WIRE": "fp8"}) == K.FP8_BLOCK
    assert K.default_wire(K.DOUBLE, "sum", {"FT_WIRE": "fp8"}) == K.DOUBLE
    import pytest
    with pytest.raises(ValueError):
        K.default_wire(K.FLOAT, "sum", {"FT_WIRE": "int4"})


def test_compiled_fast_path_module():
    """csrc/python/ftfast.c: the per-call path of Communicator.allreduce without ctypes. Loads wherever the CUDA
    library loads (no GPU needed); a null communicator is rejected by the library, bad arguments by the module."""
    import flextree._lib as L
    mod = L.fast()
    if mod is None:
        pytest.skip("lib/_ftfast.so not built (make cuda)")
    assert mod.allreduce_ex(0, 0, 0, 16, 6, 6, 0, 1.0, 0) == 1          # FT_ERR_INVALID: no communicator
    assert mod.error_peek(0) == 0
    with pytest.raises(TypeError):
        mod.allreduce_ex(0, 0)


def test_parse_topo_tree_size_suffixes():
    """'2,3+1' (a tree of N-1 plus one lonely rank) and '2,4-1' (a tree of N+1 with one phantom position): the notation of
    the reference's tree listings (cost_model/PrintTreeStructure.h), accepted by FT_TOPO."""
    assert ft.parse_topo_ex("2,3", 6) == ([2, 3], 0, 0)
    assert ft.parse_topo_ex("2,3+1", 7) == ([2, 3], 1, 0)
    assert ft.parse_topo_ex("2 4-1", 7) == ([2, 4], 0, 1)
    assert ft.parse_topo_ex("2,2,2-1", 7) == ([2, 2, 2], 0, 1)
    assert ft.parse_topo_ex("3-1", 2) == ([3], 0, 1)
    assert ft.parse_topo_ex("", 5) == ([5], 0, 0)
    assert ft.parse_topo_ex("1", 7) == ([1], 0, 0)            # ring: no tree, no suffix meaning
    for bad, n in (("2,3+1", 6), ("2,4-1", 8), ("2,3+0", 6), ("2,3-0", 6), ("2,3+9", 7)):
        with pytest.raises(ValueError):
            ft.parse_topo_ex(bad, n)
