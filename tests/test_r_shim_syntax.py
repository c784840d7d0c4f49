"""The R side (r/ate_b200_shim.c, r/ate_functions_b200.R) cannot be built or run here (no R), so it
is at least type-checked: the shim compiles against stub declarations of R's C API
(tests/r_stub/, test-only) with the real include/ate_b200.h, and every .Call in the R file names a
registered routine with the registered number of arguments."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "r", "ate_b200_shim.c")
RFILE = os.path.join(ROOT, "r", "ate_functions_b200.R")


def test_shim_type_checks_against_the_c_abi():
    gcc = shutil.which("gcc") or "/usr/bin/gcc"
    r = subprocess.run([gcc, "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-Wno-unused-parameter",
                        "-Wno-cast-function-type", "-I", os.path.join(ROOT, "tests", "r_stub"),
                        "-I", os.path.join(ROOT, "include"), SHIM], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def _registered():
    src = open(SHIM).read()
    table = dict((m.group(1), int(m.group(2)))
                 for m in re.finditer(r'\{"(C_ate_\w+)",\s*\(DL_FUNC\)&\1,\s*(\d+)\}', src))
    for name, nargs in table.items():        # the definition has that many SEXP parameters
        m = re.search(r"SEXP\s+" + name + r"\s*\(([^)]*)\)", src)
        assert m, name
        assert len([a for a in m.group(1).split(",") if a.strip() and a.strip() != "void"]) == nargs, name
    return table


def _calls(text):
    """(.Call name, number of arguments after the name) for every .Call( in the R source."""
    out = []
    for m in re.finditer(r'\.Call\("(\w+)"', text):
        i = text.index("(", m.start())
        depth, args, j, in_str = 0, 0, i, None
        while True:
            ch = text[j]
            if in_str:
                if ch == in_str:
                    in_str = None
            elif ch in "\"'":
                in_str = ch
            elif ch in "([{":
                depth += 1
            elif ch in ")]}":
                depth -= 1
                if depth == 0:
                    break
            elif ch == "," and depth == 1:
                args += 1
            elif ch == "#":                    # comment to end of line
                j = text.index("\n", j)
            j += 1
        out.append((m.group(1), args))
    return out


def test_every_dot_call_is_registered_with_matching_arity():
    table = _registered()
    assert {"C_ate_cv_glmnet", "C_ate_glm_binomial", "C_ate_lm", "C_ate_crossfit_aipw"} <= set(table)
    calls = _calls(open(RFILE).read())
    assert calls
    for name, nargs in calls:
        assert name in table, f"{name} is called from R but not registered in the shim"
        assert nargs == table[name], f"{name}: called with {nargs} arguments, registered with {table[name]}"


def test_r_bodies_keep_the_reference_entry_points():
    txt = open(RFILE).read()
    for fn in ("ate_condmean_lasso", "ate_lasso", "prop_score_lasso", "tau_hat_dr_est", "doubly_robust_glm",
               "ate_condmean_ols", "prop_score_weight", "prop_score_ols", "belloni", "crossfit_aipw"):
        assert re.search(r"^" + fn + r"\s*<-\s*function\(", txt, flags=re.M), fn
