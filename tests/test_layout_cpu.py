"""Host-side state layout of generated modules (stochaskell_b200/layout.py = flattenInput / unflattenOutput of
haskell/Language/CUDA.hs) against the CC wire format of the reference (printCC CC.hs:434-438, ccPrint
:378-396, readCC :446-456) on the soccer example of rj.ipynb cell 10."""
import numpy as np
import pytest

from stochaskell_b200.layout import flatten_input, format_cc_wire, parse_cc_wire, parse_type, unflatten_output, union_slots

SOCCER_T = "Union[(R, ZVec), (R, R, ZVec)]"


def test_soccer_wire_example_round_trips_through_the_state_fields():
    m1 = parse_cc_wire("0 2.5 5 5 1 3 2 3", SOCCER_T)                 # Model1 lambda = 2.5, y = [5,1,3,2,3]
    m2 = parse_cc_wire("1 2.5 0.0212 5 5 1 3 2 3", SOCCER_T)          # Model2 lambda kappa y
    assert m1 == (0, 2.5, [5, 1, 3, 2, 3]) and m2 == (1, 2.5, 0.0212, [5, 1, 3, 2, 3])
    f = flatten_input(0, SOCCER_T, [m1, m2])
    # tag + one slot per (kind, occurrence) shared by the constructors: r0 = lambda, r1 = kappa (unused by Model1: 0), zv0 = y
    assert sorted(f) == ["in0_r0", "in0_r1", "in0_tag", "in0_zv0", "in0_zv0_len"]
    assert f["in0_tag"].tolist() == [[0, 1]] and f["in0_r0"].tolist() == [[2.5, 2.5]] and f["in0_r1"].tolist() == [[0.0, 0.0212]]
    assert f["in0_zv0"].shape == (5, 2) and f["in0_zv0"].dtype == np.int32 and f["in0_zv0"][:, 1].tolist() == [5, 1, 3, 2, 3]
    assert f["in0_zv0_len"].tolist() == [[5, 5]]
    back = unflatten_output(0, SOCCER_T, f)
    assert back == [m1, m2]
    assert format_cc_wire(back[0], SOCCER_T) == "0 2.5 5 5 1 3 2 3"
    assert parse_cc_wire(format_cc_wire(back[1], SOCCER_T), SOCCER_T) == m2


def test_slot_sharing_follows_union_layout_of_the_lowering():
    slots, cmap = union_slots([["R", "ZVec"], ["R", "R", "ZVec"], ["Z", "RVec", "R"]])
    assert [s for s, _ in slots] == ["r0", "zv0", "r1", "z0", "rv0"]
    assert cmap == [[0, 1], [0, 2, 1], [3, 4, 0]]
    # the generated module declares the same fields
    from stochaskell_b200 import ir_to_cuda
    ir = "let v_0_0 = id i_0_0 :: Union[(R, ZVec), (R, R, ZVec)]\n in [v_0_0]\n"
    src = ir_to_cuda(ir, [SOCCER_T])
    for f in ("in0_tag", "in0_r0", "in0_r1", "in0_zv0", "in0_zv0_len"):
        assert f"field {f} state" in src


def test_scalars_vectors_and_errors():
    assert flatten_input(2, "Z", [3, 4])["in2"].tolist() == [[3, 4]]
    f = flatten_input(1, "RVec", [[1.0, 2.0, 3.0], [4.0]])
    assert f["in1"].shape == (3, 2) and f["in1_len"].tolist() == [[3, 1]] and f["in1"][:, 1].tolist() == [4.0, 0.0, 0.0]
    assert unflatten_output(1, "RVec", f) == [[1.0, 2.0, 3.0], [4.0]]
    assert parse_type("vector R") == "RVec"
    with pytest.raises(ValueError):
        parse_cc_wire("0 2.5 5 5 1 3 2 3 9", SOCCER_T)                 # readCC's `(cs,[])` pattern: every token is consumed
    with pytest.raises(ValueError):
        parse_type("RMat")
    with pytest.raises(ValueError):
        flatten_input(0, SOCCER_T, [(0, 2.5)])
