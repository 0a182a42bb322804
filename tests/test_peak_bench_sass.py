"""The roofline denominators come from vgen_b200/peak_bench (integer issue-rate microbenchmarks, run by bench.py inside the
bench run). Its numbers are "thread-level SASS instructions per second" computed from a FIXED count of 4 instructions per
unrolled step, so the count has to be what ptxas really emitted: round 1's version counted adds that ptxas had fused or moved
to the other pipe. This test disassembles the built binary (cuobjdump, no GPU needed) and checks every kernel's histogram."""
import collections
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "vgen_b200", "peak_bench")

# kernel KIND (enum order in peak_bench.cu) -> {opcode prefix: share of the 4 instructions per step}
EXPECT = {
    0: {"IMAD": 4},                                   # imad
    1: {"LOP3": 2, "SHF": 2},                         # alu_lop3_shf
    2: {"IADD3": 4},                                  # alu_iadd3
    4: {"IMAD.WIDE": 4},                              # imad_wide_carry_chain (IMAD.WIDE.U32 + IMAD.WIDE.U32.X)
    6: {"IMAD": 2, "LOP3": 1, "SHF": 1},              # mix_imad_alu_1_1
    7: {"IMAD": 1, "LOP3": 1, "SHF": 1, "IADD3": 1},  # mix_imad_alu_1_3
    8: {"IMAD.WIDE": 2, "LOP3": 1, "SHF": 1},         # mix_imadwidex_alu_1_1
}


def _histograms():
    out = subprocess.run(["cuobjdump", "-sass", BIN], capture_output=True, text=True, check=True).stdout
    hists, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : _Z1kILi(\d+)E", line)
        if m:
            cur = hists.setdefault(int(m.group(1)), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    return hists


@pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(BIN), reason="needs cuobjdump and the built peak_bench")
def test_peak_bench_kernels_execute_what_they_count():
    hists = _histograms()
    assert len(hists) == 9
    for kind, want in EXPECT.items():
        h = hists[kind]
        body = sum(h.values())
        for op, share in want.items():
            # IMAD must not swallow IMAD.WIDE / IMAD.MOV / IMAD.IADD and vice versa
            n = sum(c for k, c in h.items() if (k.startswith(op) and not (op == "IMAD" and (k.startswith("IMAD.WIDE") or k.startswith("IMAD.MOV") or k.startswith("IMAD.IADD")))))
            assert abs(n / body - share / 4) < 0.02, (kind, op, n, body, dict(h))
    # the two "as compiled" kernels: what ptxas makes of plain PTX is recorded, not assumed
    assert hists[5]["IMAD.IADD"] > 0 and hists[5]["IADD3"] > 0          # 2-input adds are spread over both pipes
    assert hists[3]["IMAD.WIDE.U32"] > 1000
