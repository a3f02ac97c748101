"""CPU: the restatement of the sum / max / min stackers (oracle/stack_oracle.c) against the
reference's own loops compiled from /root/reference (oracle/_ref, ref_simple.c)."""
import numpy as np
import pytest


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("op", ["sum", "max", "min"])
def test_port_matches_reference(orc, ref, dtype, op):
    rng = np.random.default_rng(3)
    for N, C, H, W in ((5, 1, 9, 17), (12, 3, 20, 33), (2, 1, 4, 300)):
        fr = rng.random((N, C, H, W))
        fr = (fr * 60000).astype(np.uint16) if dtype == np.uint16 else (fr * 1.2 - 0.1).astype(np.float32)
        for shifted in (False, True):
            sx = rng.integers(-6, 7, N) if shifted else None
            sy = rng.integers(-5, 6, N) if shifted else None
            for o32 in ((True, False) if (op == "sum" and dtype == np.uint16) else (True,)):
                a = orc.stack_simple(fr, op, sx, sy, o32)
                b = ref.stack_simple(fr, op, sx, sy, o32)
                assert a.dtype == b.dtype and np.array_equal(a, b)


def test_known_answers(any_oracle):
    fr = np.array([[[[1, 5, 0]]], [[[4, 2, 0]]], [[[3, 3, 9]]]], dtype=np.uint16)     # [3][1][1][3]
    assert any_oracle.stack_simple(fr, "max")[0, 0].tolist() == [4, 5, 9]
    assert any_oracle.stack_simple(fr, "min")[0, 0].tolist() == [1, 2, 0]
    assert any_oracle.stack_simple(fr, "sum", output_32bits=False)[0, 0].tolist() == [8, 10, 9]
    s32 = any_oracle.stack_simple(fr, "sum", output_32bits=True)[0, 0]
    assert s32.tolist() == [np.float32(0.8), np.float32(1.0), np.float32(0.9)]
    # x shift: frame 2 moved right by one pixel leaves its first column untouched
    out = any_oracle.stack_simple(fr, "max", shiftx=[0, 0, 1])[0, 0]
    assert out.tolist() == [4, 5, 3]
    # min skips (does not zero) pixels outside a shifted frame
    out = any_oracle.stack_simple(fr, "min", shiftx=[0, 0, 2])[0, 0]
    assert out.tolist() == [1, 2, 0]
    assert any_oracle.stack_simple(np.zeros((2, 1, 2, 2), np.uint16), "sum") is None
