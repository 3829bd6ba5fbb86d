"""The UKF kernel SOURCE (csrc/ukf.cuh) run on the CPU through tests/emu against the pinned-order C
oracle (oracle/ukf.c).  In the emulation both sides use glibc's sin/cos/atan2, so the results must
be BIT-IDENTICAL: this pins the kernel's operation order; the GPU tests then only add the
math-library difference."""
import numpy as np
import pytest

from tests import ukf_util as uu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


@pytest.mark.parametrize("L,F", [(1, 9), (4, 19), (5, 6), (8, 10), (13, 3), (16, 4), (23, 2), (32, 3)])
def test_step_bit_identical(emu, L, F):
    d = uu.random_filters(F, L, seed=L)
    want = uu.oracle_step(d)
    got = emu.emu_ukf_step(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"])
    assert np.array_equal(want[2], got[2])
    assert np.array_equal(bits(want[0]), bits(got[0])) and np.array_equal(bits(want[1]), bits(got[1]))
    assert not np.array_equal(bits(want[0]), bits(d["x"]))


def test_predict_update_halves_and_sigmas(emu):
    d = uu.random_filters(7, 8, seed=3)
    for redraw in (4, 0):
        sg_o = np.zeros((7, 7, 3))
        x1, P1, s1 = uu.cport.ukf_step(d["x"], d["P"], d["u"], None, None, flags=1 | redraw, sigmas_f=sg_o)
        x2, P2, s2 = uu.cport.ukf_step(x1, P1, None, d["z"], d["lm_pos"], d["lm_valid"], flags=2, sigmas_f=sg_o)
        sg_e = np.zeros((7, 7, 3))
        e1 = emu.emu_ukf_step(d["x"], d["P"], d["u"], None, None, flags=1 | redraw, sigmas_f=sg_e)
        assert np.array_equal(bits(sg_o), bits(sg_e))
        e2 = emu.emu_ukf_step(e1[0], e1[1], None, d["z"], d["lm_pos"], d["lm_valid"], flags=2, sigmas_f=sg_e)
        assert np.array_equal(bits(x2), bits(e2[0])) and np.array_equal(bits(P2), bits(e2[1]))
        if redraw:       # fused step == predict then update, and an update without sigmas re-draws them
            full = emu.emu_ukf_step(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], flags=7)
            assert np.array_equal(bits(full[0]), bits(e2[0])) and np.array_equal(bits(full[1]), bits(e2[1]))
            e3 = emu.emu_ukf_step(e1[0], e1[1], None, d["z"], d["lm_pos"], d["lm_valid"], flags=2)
            assert np.array_equal(bits(e3[1]), bits(e2[1]))


def test_line_record_measurements_and_match_validity(emu):
    d = uu.random_filters(11, 4, seed=5, line_records=True)
    d["lm_valid"] = np.where(d["lm_valid"] > 0, np.arange(4)[None, :], -1).astype(np.int32)   # like `match`
    want = uu.oracle_step(d, zmode=1, valid_min=0)
    got = emu.emu_ukf_step(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], valid_min=0, zmode=1)
    assert np.array_equal(bits(want[0]), bits(got[0])) and np.array_equal(bits(want[1]), bits(got[1]))


def test_no_valid_slot_is_predict_only_and_status_bits(emu):
    d = uu.random_filters(5, 4, seed=6)
    d["lm_valid"][:] = 0
    d["lm_valid"][3, 2] = 1
    d["P"][1] = -np.eye(3)                       # not positive definite: step not applied, status 1
    d["P"][4] = np.diag([0.1, 0.1, np.nan])
    want = uu.oracle_step(d)
    got = emu.emu_ukf_step(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"])
    assert list(want[2]) == [0, 1, 0, 0, 1] and np.array_equal(want[2], got[2])
    assert np.array_equal(bits(want[0]), bits(got[0])) and np.array_equal(bits(want[1]), bits(got[1]))
    pred = uu.cport.ukf_step(d["x"], d["P"], d["u"], None, None, flags=5)
    assert np.array_equal(bits(pred[0][0]), bits(got[0][0]))           # no valid slot = predict only
    assert np.array_equal(bits(got[1][1]), bits(d["P"][1]))            # failed filter left untouched
