"""CPU: the R-side .Call glue (r-package/src/melissa_R.c, shown in INTEGRATION.md section 2) cannot be built without R,
but it can be TYPE-CHECKED: against include/melissa_b200.h (every library call with the right number and types of
arguments) and against a stand-in for the R declarations it uses (tests/r_stub).  Catches a glue that has drifted from
the header, a missing include, an undeclared function."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GLUE = os.path.join(ROOT, "r-package", "src", "melissa_R.c")


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_the_call_glue_type_checks_against_the_header():
    res = subprocess.run(["gcc", "-std=gnu11", "-fsyntax-only", "-Wall", "-Werror", "-Wno-error=cast-function-type",
                          "-I", os.path.join(ROOT, "tests", "r_stub"), "-I", os.path.join(ROOT, "include"), GLUE],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr


def test_integration_md_shows_the_glue_that_is_in_the_tree():
    md = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    src = [line.rstrip() for line in open(GLUE).read().splitlines()]
    block = next(b for b in re.findall(r"```c\n(.*?)```", md, re.S) if "C_melissa_fit" in b)
    shown = [line.rstrip() for line in block.strip().splitlines() if line.strip()]
    assert len(shown) > 40
    missing = [line for line in shown if line not in src]
    assert not missing, missing[:3]


def test_the_r_wrapper_passes_the_arguments_the_glue_takes():
    # .Call(C_melissa_fit, ...) in the R wrapper: as many arguments as the C function has parameters, as many as registered
    c = open(GLUE).read()
    n_params = c[c.index("SEXP C_melissa_fit("):].split(")")[0].count("SEXP ") - 1
    registered = int(re.search(r'"C_melissa_fit", \(DL_FUNC\) &C_melissa_fit, (\d+)', c).group(1))
    r = open(os.path.join(ROOT, "r-package", "R", "melissa_inner.R")).read()
    call = r[r.index(".Call(C_melissa_fit"):]
    depth, end = 0, 0
    for i, ch in enumerate(call):
        depth += ch == "("
        depth -= ch == ")"
        if depth == 0 and i > 5:
            end = i
            break
    args, depth, cur = [], 0, ""
    for ch in call[len(".Call("):end]:
        if ch == "," and depth == 0:
            args.append(cur)
            cur = ""
            continue
        depth += ch in "(["
        depth -= ch in ")]"
        cur += ch
    args.append(cur)
    assert n_params == registered == len(args) - 1, (n_params, registered, len(args) - 1)
