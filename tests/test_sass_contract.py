"""The hardware paths the hot kernels are supposed to use, read off the built library's SASS (cuobjdump, no GPU needed): the
prefill GEMM issues tcgen05 MMAs, reads TMEM and is fed by TMA; the decode kernels run on dp4a; the persistent list kernel
streams weights with bulk copies; none of the product instantiations of the hot kernels spills.  Guards against a change
that silently drops one of them (a spill in kernel 3's epilogue costs more than any lever of DESIGN.md's table bought)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llama.cpp-q_4_0_b200", "lib", "libggml_b200.so")


@pytest.fixture(scope="module")
def kernels():
    if shutil.which("cuobjdump") is None or shutil.which("c++filt") is None:
        pytest.skip("cuobjdump / c++filt not on PATH")
    if not os.path.exists(LIB):
        pytest.skip("library not built")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    out, cur = {}, None
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            out[cur] = {}
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m and cur:
            op = m.group(1)
            for key in ("UTCIMMA", "LDTM", "UTMALDG", "UBLKCP", "IDP.4A", "FFMA2", "STL", "LDL", "SYNCS"):
                if op == key or op.startswith(key + "."):
                    out[cur][key] = out[cur].get(key, 0) + 1
    names = list(out)
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return {re.sub(r"\(.*$", "", d): out[n] for n, d in zip(names, dem)}


def _one(kernels, pattern):
    hits = [k for k in kernels if re.search(pattern, k)]
    assert hits, f"no kernel matches {pattern}"
    return [(k, kernels[k]) for k in hits]


def test_prefill_gemm_runs_on_tcgen05_tmem_and_tma(kernels):
    for pat in (r"gemm_q4_0_tc3_kernel<false, false, 4, false>", r"gemm_q4_0_tc3_kernel<false, false, 8, false>", r"gemm_q4_0_tc3_kernel<false, false, 4, true>"):
        for name, k in _one(kernels, pat):
            assert k.get("UTCIMMA", 0) >= 8, name           # data MMA + constant-tile MMA per TMEM buffer
            assert k.get("LDTM", 0) >= 8, name              # two half-fragment loads per block and ring slot
            assert k.get("UTMALDG", 0) >= 3, name           # weight tile, activation tile, activation scales
            assert k.get("FFMA2", 0) >= 128, name           # the two-FMA scale epilogue, packed
            assert k.get("STL", 0) == 0 and k.get("LDL", 0) == 0, name


def test_decode_kernels_run_on_dp4a_without_spills(kernels):
    for pat in (r"gemv_q4_0_fused_kernel<4, 1, 1024, false>", r"gemv_q4_0_neigh_kernel<4, 1024, 1>", r"gemv_q4_0_fused_p2p_kernel<4, 1024>"):
        for name, k in _one(kernels, pat):
            assert k.get("IDP.4A", 0) >= 32, name
            assert k.get("STL", 0) == 0 and k.get("LDL", 0) == 0, name
    for name, k in _one(kernels, r"gemv_q4_0_neigh_kernel<4, 512, 2>"):
        assert k.get("IDP.4A", 0) >= 64, name               # two weight streams


def test_persistent_list_kernel_streams_with_bulk_copies(kernels):
    for name, k in _one(kernels, r"gemv_q4_0_set_kernel<12, false, true, false, false>"):
        assert k.get("UBLKCP", 0) >= 1 and k.get("IDP.4A", 0) >= 32 and k.get("SYNCS", 0) >= 4, name
        assert k.get("STL", 0) == 0 and k.get("LDL", 0) == 0, name
