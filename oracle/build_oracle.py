"""Build recipe for the checker libraries.  TEST INFRASTRUCTURE ONLY.

  oracle/libchronos_oracle.so          <- oracle/chronos_oracle.c (our CPU restatement)
  oracle/_ref/libchronos_ref_filter.so <- /root/reference/chronos_sensor.py:28-47 (the
                                          reference's own C helpers, read in place) +
                                          oracle/ref_filter_wrap.c

The second one is only (re)built where /root/reference exists (the build container);
the GPU box uses the prebuilt file that travels with the snapshot.  Outputs of the
reference build go to oracle/_ref/ only, which is git-ignored.
"""
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SENSOR = "/root/reference/chronos_sensor.py"
ORACLE_SO = os.path.join(HERE, "libchronos_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libchronos_ref_filter.so")
# ends_with + starts_with as they sit in the reference's BPF program text
REF_HELPER_LINES = (28, 47)


def _newer(target, *sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def build_oracle(force=False):
    src = os.path.join(HERE, "chronos_oracle.c")
    if not force and _newer(ORACLE_SO, src):
        return ORACLE_SO
    subprocess.check_call(
        ["gcc", "-O2", "-std=c11", "-Wall", "-Wextra", "-fPIC", "-shared", "-o", ORACLE_SO, src]
    )
    return ORACLE_SO


def build_ref(force=False):
    """Compile the reference's own helpers. Returns the .so path or None if neither the
    reference nor a prebuilt library is available."""
    wrap = os.path.join(HERE, "ref_filter_wrap.c")
    if not os.path.exists(REF_SENSOR):
        return REF_SO if os.path.exists(REF_SO) else None
    if not force and _newer(REF_SO, wrap, REF_SENSOR):
        return REF_SO
    os.makedirs(REF_DIR, exist_ok=True)
    with open(REF_SENSOR, "r", encoding="utf-8") as f:
        lines = f.read().split("\n")
    lo, hi = REF_HELPER_LINES
    helpers = "\n".join(lines[lo - 1:hi])
    assert "ends_with" in helpers and "starts_with" in helpers, "reference layout changed"
    with tempfile.TemporaryDirectory(prefix="chronos_ref_") as td:
        tu = os.path.join(td, "ref_filter_tu.c")
        with open(tu, "w") as f:
            f.write(helpers + "\n")
            f.write('#line 1 "ref_filter_wrap.c"\n')
            f.write(open(wrap).read())
        subprocess.check_call(
            ["gcc", "-O2", "-std=gnu11", "-w", "-fPIC", "-shared", "-o", REF_SO, tu]
        )
    return REF_SO


if __name__ == "__main__":
    print(build_oracle(force=True))
    print(build_ref(force=True))
