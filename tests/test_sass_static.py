"""Static checks on the built library's SASS (no GPU needed): the hot loops keep the instruction mix DESIGN.md section 4
argues from -- packed FP32x2 arithmetic, the MUFU count per evaluation, no register moves or scalar FP32 in the loop --
and the light list is staged by TMA bulk copies. A compiler or source change that silently breaks one of these shows up
here before anyone spends GPU time on it."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "volumerenderer_b200", "libvr.so")

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(LIB),
                                reason="needs cuobjdump and a built libvr.so")


def _functions():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs = {}
    name = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = []
            continue
        m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m and name:
            funcs[name].append((int(m.group(1), 16), m.group(2)))
    return funcs


@pytest.fixture(scope="module")
def funcs():
    return _functions()


def _inner_mufu_loop(ins):
    """the shortest backward-branch loop that contains MUFU: the per-light-pair body"""
    loops = []
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            body = [x for b, x in ins if int(m.group(1), 16) <= b <= a]
            if any("MUFU" in x for x in body):
                loops.append(body)
    assert loops, "no loop with MUFU found"
    return min(loops, key=len)


def _count(body, pat):
    return sum(bool(re.search(pat, t)) for t in body)


def _find(funcs, *parts):
    hits = [n for n in funcs if all(p in n for p in parts)]
    assert hits, parts
    return funcs[hits[0]]


@pytest.mark.parametrize("mangled, evals, packed_max, mufu", [
    # (threads, CTAs/SM, sphere, entry length, slice units)
    ("k_eval_segILi256ELi2ELb0ELi16ELb0", 64, 152, 32),   # point, 16-sample entries, full units: 4.75 FMA-pipe cycles, 0.5 MUFU per evaluation
    ("k_eval_segILi256ELi2ELb0ELi16ELb1", 64, 152, 32),   # ... slice units (the last wave)
    ("k_eval_segILi256ELi2ELb0ELi8ELb1", 32, 88, 16),     # point, 8-sample entries (odd segments, all as slice units): 5.5 cycles
    ("k_eval_segILi256ELi2ELb1ELi8ELb0", 32, 120, 48),    # sphere: 1.5 MUFU per evaluation
    ("k_eval_segILi256ELi2ELb1ELi8ELb1", 32, 120, 48),
])
def test_segment_kernel_hot_loop_mix(funcs, mangled, evals, packed_max, mufu):
    body = _inner_mufu_loop(_find(funcs, mangled))
    packed = _count(body, r"FFMA2|FMUL2|FADD2")
    assert packed <= packed_max, (packed, packed_max)
    assert packed >= 0.9 * packed_max            # the loop found is the pair body, not a tail
    assert _count(body, "MUFU") == mufu
    assert _count(body, r"(^|\s)MOV |IMAD\.MOV") == 0, "register moves in the hot loop"
    assert _count(body, r"\bFFMA |\bFMUL |\bFADD ") == 0, "scalar FP32 in the packed loop"
    assert _count(body, r"LDS\.128") == 4        # two lights, two broadcast loads each
    assert _count(body, r"STL|LDL") == 0, "spills in the hot loop"


def _mufu_use_distances(body):
    """Instructions between every MUFU of the loop body and the first instruction that reads its result."""
    out = []
    for i, t in enumerate(body):
        m = re.match(r"MUFU\.\w+ (R\d+),", t)
        if not m:
            continue
        r = m.group(1)
        n = int(r[1:])
        pair = "R%d" % (n - n % 2)
        for j in range(i + 1, len(body)):
            ops = body[j].split(" ", 1)[1] if " " in body[j] else ""
            srcs = ops.split(",")[1:]
            if any(re.search(r"\b%s\b" % r, s_) or ("F32x2" in s_ and re.search(r"\b%s\b" % pair, s_)) for s_ in srcs):
                out.append(j - i)
                break
    return out


@pytest.mark.parametrize("mangled", ["k_eval_segILi256ELi2ELb0ELi16ELb0", "k_eval_segILi256ELi2ELb0ELi16ELb1"])
def test_segment_kernel_schedule_hides_the_reciprocal_latency(funcs, mangled):
    """Same instruction mix, 5 % slower: with one register too many ptxas serialises the samples of an entry (every
    reciprocal consumed 2-5 instructions after its MUFU; measured 10.28 vs 9.76 ms on C2, `wait` stalls 2.0 vs 1.1 per
    issue). The full units keep their bookkeeping uniform (chunk = blockIdx.x) so that the loop has the registers to
    interleave two or three samples: the typical reciprocal is consumed ten or more instructions later."""
    d = sorted(_mufu_use_distances(_inner_mufu_loop(_find(funcs, mangled))))
    assert len(d) == 32
    assert d[len(d) // 2] >= 9, d


def test_exact_list_kernel_hot_loop(funcs):
    body = _inner_mufu_loop(_find(funcs, "k_eval_point3ILi8ELi3ELi2ELb1"))
    assert _count(body, r"FFMA2|FMUL2|FADD2") == 56 and _count(body, "MUFU") == 16   # 16 evaluations of 8 samples x 2 lights
    assert _count(body, r"(^|\s)MOV |IMAD\.MOV") == 0 and _count(body, r"STL|LDL") == 0


def test_light_list_is_staged_by_tma(funcs):
    for part in ("k_eval_segILi256ELi2ELb0ELi16ELb0", "k_eval_segILi256ELi2ELb0ELi16ELb1", "k_eval_segILi256ELi2ELb1ELi8ELb0",
                 "k_eval_point3ILi8ELi3ELi2ELb0", "k_eval_sphere2ILi8ELb0"):
        ins = _find(funcs, part)
        assert any("UBLKCP" in t for _, t in ins), part          # cp.async.bulk global -> shared
        assert any("SYNCS" in t for _, t in ins), part           # mbarrier arrive / try_wait
