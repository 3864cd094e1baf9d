"""GPU: the reference's OWN acceptance callers linked against libdefl_b200.so.

`make -C oracle b200tests` (run by __graft_entry__.build() in the container, where /root/reference exists)
compiles /root/reference/test/test_files.c and test_fuzz.c UNMODIFIED against include/defl/*.h and links them with
defl_b200/libdefl_b200.so; the binaries and a copy of the reference's test data live under the git-ignored
oracle/_ref/ and travel to the GPU box.  Nothing here reads /root/reference at run time.

test_files.c: corpus one-shot (test_files.c:174-257), chunk lists 1/2/4/8 bytes (:260-331), streaming in
64..1024-byte pieces (:688-790), byte-wise streaming with empty polls (:792-869), error inputs (:460-544),
regressions (:617-686), streaming edge cases (:871-1029); the three huff unit checks (:546-615) run against the
stand-in oracle/huff/huff.h, as they do for the reference build.  Expected: the same 176 checks, 0 failed.
test_fuzz.c:138-311: return codes in {OK, ERR, EFULL}, no crash; and the same number of successes as the
reference build of the same fuzzer (it is seeded)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFOUT = os.path.join(ROOT, "oracle", "_ref")
ANSI = re.compile(r"\x1b\[[0-9;]*m")


def _run(binary, args, cwd, timeout):
    exe = os.path.join(REFOUT, binary)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (make -C oracle b200tests needs /root/reference)")
    p = subprocess.run([exe] + args, cwd=cwd, capture_output=True, text=True, timeout=timeout)
    return p.returncode, ANSI.sub("", p.stdout + p.stderr)


@pytest.mark.gpu
def test_reference_test_files_against_b200_library():
    rc, out = _run("test_files_b200", [], os.path.join(REFOUT, "test"), 600)
    m = re.search(r"(\d+) tests ran, (\d+) passed, (\d+) failed", out)
    assert m, out[-2000:]
    ran, passed, failed = map(int, m.groups())
    failures = [l for l in out.splitlines() if "✘" in l or "FAIL" in l.upper() and "0 failed" not in l]
    assert failed == 0 and rc == 0, "\n".join(failures[:40]) + out[-1500:]
    assert ran == 176 and passed == 176, (ran, passed)
    t = re.search(r"test results \(([\d.]+)s\)", out)
    assert t and float(t.group(1)) < 120.0, out[-500:]


@pytest.mark.gpu
def test_reference_fuzzer_against_b200_library():
    n = 10000
    rc, out = _run("test_fuzz_b200", [str(n)], REFOUT, 900)
    m = re.search(r"Successful: (\d+).*?Expected errors: (\d+).*?Crashes/Issues: (\d+)", out, re.S)
    assert m and rc == 0, out[-1500:]
    ok, err, bad = map(int, m.groups())
    assert bad == 0 and ok + err == n
    rrc, rout = _run("test_fuzz", [str(n)], REFOUT, 900)          # the reference's own build of the same fuzzer
    rm = re.search(r"Successful: (\d+).*?Expected errors: (\d+).*?Crashes/Issues: (\d+)", rout, re.S)
    assert rm and (int(rm.group(1)), int(rm.group(2))) == (ok, err), (rm.groups() if rm else None, (ok, err))
