"""What the compiled sm_100a code contains, checked here without a GPU (cuobjdump -sass):
the TMA tile load and mbarrier in the hashing kernel, the ALU/FMA pipe split that the
roofline argument in DESIGN.md rests on, and the instruction mix the pipe probes claim."""
import collections
import re
import shutil
import subprocess

import pytest

import __graft_entry__ as entry
from bee_code_interpreter_fs_b200 import _build

CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"


@pytest.fixture(scope="module")
def functions():
    if not shutil.which(CUOBJDUMP):
        pytest.skip("cuobjdump not available")
    entry.build()
    text = subprocess.run([CUOBJDUMP, "-sass", _build.LIB_PATH], capture_output=True, text=True, check=True).stdout
    out = {}
    for chunk in re.split(r"\n\s*Function : ", text)[1:]:
        name, body = chunk.split("\n", 1)
        ins = []
        for line in body.split("\n"):
            m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", line)
            if m:
                ins.append((int(m.group(1), 16), re.sub(r"^@!?U?P\d+\s+", "", m.group(2).strip())))
        out[name.strip()] = ins
    return out


def hot_loop(ins):
    """Instructions of the largest backward-branch region (the per-block loop)."""
    best = None
    for addr, text in ins:
        m = re.search(r"BRA(?:\.U)?(?:\.ANY)?\s+(?:!?U?P\d,\s*)?0x([0-9a-f]+)", text)
        if m and int(m.group(1), 16) < addr:
            lo = int(m.group(1), 16)
            if best is None or addr - lo > best[1] - best[0]:
                best = (lo, addr)
    return [t for a, t in ins if best[0] <= a <= best[1]]


def mix(texts):
    c = collections.Counter()
    for t in texts:
        op = t.split()[0]
        c["IMAD.HI" if op.startswith("IMAD.HI") else "IMAD.WIDE" if op.startswith("IMAD.WIDE") else op.split(".")[0]] += 1
    return c


def find(functions, pattern):
    names = [n for n in functions if re.search(pattern, n)]
    assert names, f"no kernel matches {pattern}"
    return functions[names[0]]


def test_code_is_sm100_native_and_uses_tma_and_mbarrier(functions):
    loop = hot_loop(find(functions, r"kernel_tmaILi3ELi2ELi256ELi2ELi1E"))
    ops = mix(loop)
    assert ops["UTMALDG"] == 1, "one TMA tile load per block index"
    assert ops["SYNCS"] >= 2, "mbarrier arrive + try_wait"
    assert ops["LDS"] == 4 and ops["LDG"] == 0, "the hot loop reads shared memory only"
    assert not any(op in ops for op in ("HMMA", "IMMA", "LDL", "STL")), ops
    bulk = mix(t for _, t in find(functions, r"kernel_bulkILi3"))
    assert bulk["UBLKCP"] >= 1


def test_adds_run_on_the_fma_pipe_and_the_alu_count_is_near_the_floor(functions):
    """FIPS work per block: 672 shifts/rotates + 352 three-input booleans + 16 byte swaps stay on
    the ALU pipe (1 040); at FMA level 3 the ~600 adds are IMADs, at level 0 most are IADD3."""
    l3 = mix(hot_loop(find(functions, r"kernel_tmaILi3ELi2ELi256ELi2ELi1E")))
    l0 = mix(hot_loop(find(functions, r"kernel_tmaILi0ELi2ELi256ELi2ELi1E")))
    assert 672 <= l3["SHF"] <= 680 and 352 <= l3["LOP3"] <= 380 and l3["PRMT"] == 16
    assert 590 <= l3["IMAD"] <= 640 and l3["IADD3"] <= 8
    assert l0["IADD3"] >= 200 and l0["IMAD"] <= 200
    alu3 = sum(l3[k] for k in ("SHF", "LOP3", "PRMT", "IADD3", "VIADD", "ISETP", "SEL", "PLOP3", "LEA", "MOV"))
    assert alu3 <= 1130, alu3  # includes ~50 instructions of the once-per-message padding branch
    l4 = mix(hot_loop(find(functions, r"kernel_tmaILi4ELi2ELi256ELi2ELi1E")))
    assert l4["IMAD.HI"] == 96 and 576 <= l4["SHF"] <= 584


def test_pipe_probe_instruction_mix_matches_its_claim(functions):
    want = {0: {"SHF": 64, "LOP3": 64, "IADD3": 64}, 2: {"IMAD": 64}, 3: {"IMAD.HI": 64},
            4: {"SHF": 64, "LOP3": 64, "IADD3": 64, "IMAD": 128}, 6: {"SHF": 64, "LOP3": 64, "IMAD": 128}}
    for mode, ops in want.items():
        got = mix(hot_loop(find(functions, rf"kernel_peakILi{mode}E")))
        for op, count in ops.items():
            assert got[op] == count, (mode, op, got)
        if mode != 3:  # IMAD.HI wants register pairs: ptxas adds moves around it
            assert sum(got.values()) <= sum(ops.values()) + 4  # loop counter and branch only


def test_warp_per_object_kernel_round_chain_is_short(functions):
    """kernel_long<FMA 3, adds 4, M>: phase B's loop (the serial chain) keeps 11 ALU-pipe
    instructions per round -- 6 SHF + 4 LOP3 + 1 IADD3 -- and runs the other adds as IMAD."""
    for m in (1, 2, 4, 8, 16, 32):
        ins = find(functions, rf"kernel_longILi3ELi4ELi{m}ELi32ELi0E")
        ops = mix(t for _, t in ins)
        assert ops["LDS"] >= 16 and ops["STS"] >= 16
        assert ops["SHF"] >= 384 + 288  # 64 rounds x 6 in phase B + 48 schedule steps x 6 in phase A
    ins = find(functions, r"kernel_longILi3ELi4ELi1ELi32ELi0E")
    lds = [a for a, t in ins if t.startswith("LDS")]
    back = [(a, int(re.search(r"0x([0-9a-f]+)", t).group(1), 16)) for a, t in ins
            if t.startswith("BRA") and re.search(r"0x([0-9a-f]+)", t) and int(re.search(r"0x([0-9a-f]+)", t).group(1), 16) < a]
    end, start = min((a, lo) for a, lo in back if a > lds[-1])  # the loop that closes after the last LDS
    chain = mix(t for a, t in ins if start <= a <= end)
    assert chain["SHF"] == 384 and chain["LOP3"] == 256 and chain["LDS"] == 16
    assert chain["IADD3"] + chain["VIADD"] <= 80 and 300 <= chain["IMAD"] <= 340
