"""The row loop of the whole-solve kernel, checked in the machine code of the built library (no GPU): instruction count per
row within budget, no local-memory traffic, next to no special-register reads -- see tools/row_loop_check.py and DESIGN.md
section 3b for what happens to wide rows when this loop grows unnoticed."""
import importlib.util
import os
import shutil

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tool():
    spec = importlib.util.spec_from_file_location("row_loop_check", os.path.join(ROOT, "tools", "row_loop_check.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_row_loop_of_every_whole_solve_instance_is_within_budget():
    tool = _tool()
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    if not os.path.exists(tool.DEFAULT_LIB):
        pytest.skip("libsart_b200.so not built")
    rows, bad = tool.check()
    solve = [r for r in rows if r[0] == "solve"]
    assert len(solve) == 16, f"expected 16 instances of fused_solve_kernel, found {len(solve)}"
    assert not bad, "\n".join(bad)


def test_batch_ring_kernels_use_tensor_copies_and_keep_their_stage_loops_in_registers():
    """The two ring kernels of sart_solve_batch (csrc/batch_ring.cuh): TMA tensor copies (UTMALDG) feed them, and the loop
    that multiplies a stage out of shared memory touches no local memory."""
    tool = _tool()
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    if not os.path.exists(tool.DEFAULT_LIB):
        pytest.skip("libsart_b200.so not built")
    found = 0
    for name, ins in tool.functions(tool.DEFAULT_LIB).items():
        if "batch_ring_kernel" not in name:
            continue
        found += 1
        ops = [tool.opcode(t) for _, t in ins]
        assert "UTMALDG" in ops, name
        n, c = tool.row_loop(ins, 128)  # a stage's FMAs: 256 (forward), 512 (back)
        assert c["STL"] + c["LDL"] == 0 and c["LDS"] >= 16 and c["LDG"] == 0, (name, n, dict(c))
    assert found == 2
