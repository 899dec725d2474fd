#!/usr/bin/env python
"""Size of the row loop of every fused_solve_kernel instance in libsart_b200.so, read from its SASS (no GPU needed).

    python tools/row_loop_check.py [library] [--all]

The whole-solve kernel sits at its 128-register cap. When the code around the row loop grows, ptxas starts to re-derive loop
invariants inside that loop (S2R of thread / cluster ids, parameter loads, the multiplies behind shared-memory addresses) or
to spill there; at a cluster of 4 the sweep is HBM-bound and hides that, wide rows (112 SMs streaming) lose it one to one
(287 instead of 239 instructions per row cost 16 % at 200 000 voxels). The row loop is taken to be
the smallest loop (backward branch) that holds the row's 8 K FMAs. Exit code 1 when an instance exceeds the budget
165 + 12.5 K instructions, touches local memory (STL / LDL) or reads more than four special registers per row.
`tests/test_row_loop_budget.py` runs this as part of the CPU suite."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_LIB = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "libsart_b200.so")


def functions(lib):
    """{mangled name: [(address, instruction text)]} of every function in the library."""
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    out, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = out.setdefault(m.group(1), [])
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur is not None:
            cur.append((int(m.group(1), 16), m.group(2).strip()))
    return out


def opcode(text):
    parts = text.split()
    return (parts[1] if parts[0].startswith("@") else parts[0]).split(".")[0]


def row_loop(ins, min_ffma):
    """Counter of opcodes of the smallest loop with at least min_ffma FFMA, or None."""
    index = {a: i for i, (a, _) in enumerate(ins)}
    best = None
    for i, (a, text) in enumerate(ins):
        m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", text)
        if not m:
            continue
        target = int(m.group(1), 16)
        if target > a or target not in index:
            continue
        body = ins[index[target]:i + 1]
        c = collections.Counter(opcode(t) for _, t in body)
        if c["FFMA"] >= min_ffma and (best is None or len(body) < best[0]):
            best = (len(body), c)
    return best


def check(lib=DEFAULT_LIB, every=False):
    rows, bad = [], []
    for name, ins in functions(lib).items():
        m = re.search(r"fused_(solve|sweep)_kernelILi(\d+)ELi(\d+)ELi(\d+)", name)
        if not m or (m.group(1) == "sweep" and not every):
            continue
        kind, k, d = m.group(1), int(m.group(3)), int(m.group(4))
        loop = row_loop(ins, 8 * k)
        if loop is None:
            bad.append(f"{kind} K={k} D={d}: no row loop found")
            continue
        n, c = loop
        special = c["S2R"] + c["CS2R"] + c["S2UR"]
        local = c["STL"] + c["LDL"]
        budget = 165 + 12.5 * k
        rows.append((kind, d, k, n, budget, c["FFMA"], c["LDS"], special, c["LDC"] + c["LDCU"] + c["ULDC"], local, len(ins)))
        if kind == "solve" and (n > budget or local or special > 4):
            bad.append(f"solve K={k} D={d}: {n} instructions per row (budget {budget:.0f}), {special} special-register reads, "
                       f"{local} local-memory accesses")
    rows.sort()
    return rows, bad


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    rows, bad = check(args[0] if args else DEFAULT_LIB, "--all" in sys.argv)
    print(f"{'kernel':<7}{'D':>3}{'K':>3}{'row loop':>10}{'budget':>8}{'FFMA':>6}{'LDS':>5}{'S2R':>5}{'LDC':>5}{'local':>7}{'total':>7}")
    for kind, d, k, n, budget, ffma, lds, special, ldc, local, total in rows:
        print(f"{kind:<7}{d:>3}{k:>3}{n:>10}{budget:>8.0f}{ffma:>6}{lds:>5}{special:>5}{ldc:>5}{local:>7}{total:>7}")
    for line in bad:
        print("OVER BUDGET: " + line)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
