"""TEST INFRASTRUCTURE ONLY -- materialises the reference's own hot-path Python files under oracle/_ref/.

The reference (hrukalive/DiffSinger) is pure Python, so "building" it means making the files it imports for this path
available where /root/reference does not exist (the GPU box): this recipe imports the unmodified reference from
/root/reference (oracle/ref_harness.py), asks the interpreter which files of that tree the path actually loaded
(modules.core / modules.backbones / the two vendored solvers, plus the acoustic inference pipeline and top-level model that
BASELINE config #1 runs through), and copies exactly those, byte for byte, to oracle/_ref/ together with the data files
config #1 reads (samples/00_*.ds, the dictionary, the YAML configs).  oracle/_ref/ is git-ignored (never part of the
history) but NOT gpurun-ignored, so it travels to the GPU box like the built .so.  `bench.py --impl reference` and the
`cpu_baseline` leg import the reference from there (`cpu_baseline.kind: "reference"`); when it is absent they fall back to
the oracle port and say so.

Run by `__graft_entry__.build()` when /root/reference is present; a no-op elsewhere.
"""
from __future__ import annotations

import glob
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SRC = "/root/reference"
DATA = ["samples/00_*.ds", "dictionaries/opencpop-extension.txt", "configs/base.yaml", "configs/acoustic.yaml",
        "configs/variance.yaml"]


def build_ref(dest: str = DEST) -> bool:
    if not os.path.isfile(os.path.join(SRC, "modules", "backbones", "wavenet.py")):
        return os.path.isdir(dest)
    os.environ["DS_REFERENCE_ROOT"] = SRC
    root = os.path.dirname(HERE)
    if root not in sys.path:
        sys.path.insert(0, root)
    from oracle.ref_harness import load_reference
    load_reference()
    import inference.dpm_solver_pytorch  # noqa: F401  (imported lazily by ddpm.inference, ddpm.py:247,286)
    import inference.uni_pc  # noqa: F401
    import inference.ds_acoustic  # noqa: F401  (config #1: .ds -> tensors -> DiffSingerAcoustic)
    import modules.toplevel  # noqa: F401
    files = sorted({os.path.relpath(m.__file__, SRC) for m in list(sys.modules.values())
                    if getattr(m, "__file__", None) and os.path.abspath(m.__file__).startswith(SRC + os.sep)})
    for pat in DATA:
        files += [os.path.relpath(p, SRC) for p in sorted(glob.glob(os.path.join(SRC, pat)))]
    if os.path.isdir(dest):
        shutil.rmtree(dest)
    manifest = {}
    for rel in files:
        out = os.path.join(dest, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), out)
        manifest[rel] = hashlib.sha256(open(out, "rb").read()).hexdigest()[:16]
    with open(os.path.join(dest, "MANIFEST.json"), "w") as f:
        json.dump({"source": SRC, "files": manifest}, f, indent=1)
    return True


if __name__ == "__main__":
    ok = build_ref()
    print("oracle/_ref", "ready" if ok else "unavailable (no /root/reference and no earlier copy)")
