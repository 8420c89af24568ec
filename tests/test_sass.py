"""Structure of the shipped SASS (no GPU needed: cuobjdump reads the cross-compiled library).

Guards the properties DESIGN.md section 4 argues from: the recall scorer issues tcgen05.mma for CTA pairs from
a lean loop (operands warp-uniform, ONE elect per K block -- a loop that lost that had 6 ELECT and 22 R2UR per
K block and cost 6 % of the headline), the seed chain carries griddepcontrol, TMA loads carry no multicast."""
import re
import shutil
import subprocess
from pathlib import Path

import pytest

LIB = Path(__file__).resolve().parent.parent / "omnia_b200" / "lib" / "libomnia_recall.so"

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None or not LIB.exists(),
                                reason="needs cuobjdump and the built library")


@pytest.fixture(scope="module")
def kernels():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    out, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            out[cur] = []
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
        if m and cur:
            out[cur].append(m.group(1).strip())
    return out


def _one(kernels, *needles):
    hits = [k for k in kernels if all(n in k for n in needles)]
    assert len(hits) == 1, (needles, hits)
    return kernels[hits[0]]


def _ops(body, name):
    return [i for i, ins in enumerate(body) if re.search(r"(^|\s)" + re.escape(name) + r"(\.|\s|$)", ins)]


@pytest.mark.parametrize("needles,mma", [(("score_umma2_kernel", "ILi2ELi256"), "UTCHMMA.2CTA"),
                                         (("score_umma2_kernel", "ILi1ELi256"), "UTCHMMA.2CTA"),
                                         (("score_umma_kernel", "ILi2ELi256"), "UTCHMMA"),
                                         (("score_umma_kernel", "ILi0ELi32"), "UTCHMMA")])
def test_mma_issue_loop_is_lean(kernels, needles, mma):
    body = _one(kernels, *needles)
    at = [i for i, ins in enumerate(body) if ins.split()[0 if not ins.startswith("@") else 1].startswith(mma.split(".")[0])
          and (".2CTA" in ins) == (".2CTA" in mma)]
    assert len(at) == 4, "one K block = four tcgen05.mma of K=16"
    assert at[-1] - at[0] == 3, "the four MMAs of a K block are issued back to back"
    window = body[max(0, at[0] - 60):at[-1] + 8]
    assert sum("ELECT" in w for w in window) <= 1, "operands are not warp-uniform any more: ELECT waterfall around the MMAs"
    assert not any("R2UR.BROADCAST" in w for w in window)


def test_seed_chain_uses_dependent_launch(kernels):
    # griddepcontrol.launch_dependents -> PREEXIT, griddepcontrol.wait -> ACQBULK
    assert any("PREEXIT" in i for i in _one(kernels, "score_umma_kernel", "ILi0ELi32"))
    seed = _one(kernels, "emit_seed_kernel")
    assert any("PREEXIT" in i for i in seed) and any("ACQBULK" in i for i in seed)
    for needles in (("score_umma2_kernel", "ILi2ELi256"), ("score_umma_kernel", "ILi2ELi256")):
        body = _one(kernels, *needles)
        assert sum("ACQBULK" in i for i in body) >= 3, "epilogue, drain and refresh warps each wait for the seed"


def test_no_tma_multicast_load_anywhere(kernels):
    # TMA multicast loads here read only this GPU's memory, and the library has none at all; commits may multicast
    for name, body in kernels.items():
        for ins in body:
            if "UTMALDG" in ins or "UBLKCP" in ins:
                assert "MULTICAST" not in ins.upper(), (name, ins)


def test_emission_scorer_fits_beside_a_post_block(kernels):
    res = subprocess.run(["cuobjdump", "-res-usage", str(LIB)], capture_output=True, text=True, check=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+)", line)
        if m and cur:
            regs[cur] = int(m.group(1))
    scorer = [v for k, v in regs.items() if "score_umma2_kernel" in k and "ILi2ELi256" in k]
    post = [v for k, v in regs.items() if "emit_post_kernel" in k]
    assert scorer and post
    # 256 threads each, registers allocated in units of 8 per thread: both must fit the 64k-register file together
    up8 = lambda r: (r + 7) // 8 * 8
    assert (up8(scorer[0]) + up8(post[0])) * 256 <= 65536
