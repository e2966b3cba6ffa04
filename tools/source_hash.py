"""sha256 over the CUDA library's sources (csrc/ without plugin/ and generated json, include/pantr_b200.h), sorted by
path: identifies the kernels independently of a rebuild of the .so.  Used by tools/make_traffic.py and bench.py."""
import hashlib
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def source_sha256():
    csrc = os.path.join(ROOT, "pantr_cdc2023_experiments_b200", "csrc")
    files = sorted(os.path.join(dp, f) for dp, _, fs in os.walk(csrc) for f in fs if "plugin" not in dp and not f.endswith(".json"))
    files.append(os.path.join(ROOT, "include", "pantr_b200.h"))
    h = hashlib.sha256()
    for f in files:
        h.update(os.path.relpath(f, ROOT).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()


if __name__ == "__main__":
    print(source_sha256())
