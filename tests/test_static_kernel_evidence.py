"""CPU: static evidence about the compiled kernels (cuobjdump / ptxas output; no GPU needed).
The design claims checked here: the hot loop is packed FP32 (FFMA2/FADD2/FMUL2) + MUFU.RSQ, tensor
cores and TMA are not used (north_star: the path is FMA/SFU-bound), nothing spills, one cubin for
sm_100a."""
import os
import re
import subprocess

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def sass(nb):
    r = subprocess.run(["cuobjdump", "-sass", nb.LIB_PATH], capture_output=True, text=True)
    if r.returncode != 0 or not r.stdout:
        pytest.skip("cuobjdump not available")
    return r.stdout


def function_bodies(sass):
    return {f.split("\n")[0].strip(): f for f in re.split(r"\n\s*Function : ", sass)[1:]}


def test_fast_kernels_are_packed_fp32(sass):
    bodies = function_bodies(sass)
    fast = {k: v for k, v in bodies.items() if "force_fast_kernel" in k}
    assert len(fast) == 8                                    # I in {1,2,4,7} x {uniform mass, general}
    for name, body in fast.items():
        ops = re.findall(r"/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", body)
        assert ops.count("FFMA2") > 0 and ops.count("FADD2") > 0 and ops.count("FMUL2") > 0, name
        assert "MUFU" in ops, name
        scalar = sum(ops.count(o) for o in ("FFMA", "FMUL"))
        assert scalar < 0.1 * ops.count("FFMA2"), (name, scalar)      # the arithmetic is packed, not scalar


def test_no_tensor_core_or_tma_instructions(sass):
    for mnemonic in ("HMMA", "UTCHMMA", "UTCQMMA", "IMMA", "DMMA", "HGMMA", "UTMALDG", "UBLKCP", "LDTM", "STTM"):
        assert not re.search(r"\s" + mnemonic + r"\b", sass), mnemonic


def test_strict_kernel_has_no_contracted_products_of_its_own(sass):
    """The strict kernel's own multiplies and adds must stay separate instructions; the FFMAs present
    belong to the IEEE sqrt/divide expansions."""
    body = [v for k, v in function_bodies(sass).items() if "force_strict_kernel" in k][0]
    ops = re.findall(r"/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", body)
    assert sum(o.startswith("FMUL") for o in ops) >= 12 and sum(o.startswith("FADD") for o in ops) >= 9
    assert any(o.startswith("MUFU.RSQ") for o in ops) and any(o.startswith("MUFU.RCP") for o in ops)
    assert not any(o.startswith(("FFMA2", "FMUL2", "FADD2")) for o in ops)


def test_hot_kernels_do_not_spill():
    log = os.path.join(ROOT, "n-body-simulation_b200", "csrc", "ptxas.log")
    if not os.path.exists(log):
        pytest.skip("ptxas.log not present (library not built by this checkout's Makefile)")
    text = open(log).read()
    found = 0
    for m in re.finditer(r"Function properties for (\S+)\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads"
                         r"\nptxas info\s+: Used (\d+) registers", text):
        name, _, st, ld, regs = m.groups()
        if re.search(r"force_|finish_kernel|aos_|energy_kernel", name):
            found += 1
            assert st == "0" and ld == "0", name            # the one-off generator kernel may keep its RNG state on the stack
            assert int(regs) <= 255
    assert found >= 12
