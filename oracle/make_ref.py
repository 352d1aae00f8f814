"""TEST INFRASTRUCTURE ONLY. Recipe for oracle/_ref/: the UNMODIFIED reference sources of the hot path, copied from the
reference mount so that they travel to the GPU box (which has no /root/reference) and `bench.py --impl reference` /
`cpu_baseline` can time the reference's own code (cpu_baseline.kind = "reference") instead of the oracle port.

    python oracle/make_ref.py          (also run by __graft_entry__.build() when /root/reference is present)

oracle/_ref/ is git-ignored (never committed: the repo holds no reference sources) but not gpurun-ignored.
Copied (Python files only):
  third_party/act/policy.py, third_party/act/detr/**        -- ACTPolicy / DETRVAE / backbone / transformer
  robo_manip_baselines/**                                   -- RolloutAct.infer_policy, RolloutBase.get_state / get_images,
                                                               DataUtils.normalize_data / denormalize_data
The reference is pure Python: there is nothing to compile.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("RMB_REFERENCE_SRC", "/root/reference")
DEST = os.path.join(HERE, "_ref")
TREES = ("third_party/act/detr", "robo_manip_baselines")
FILES = ("third_party/act/policy.py",)


def make_ref(verbose=True):
    if not os.path.isdir(REF_SRC):
        if verbose:
            print(f"[make_ref] {REF_SRC} not present: keeping whatever oracle/_ref already holds")
        return os.path.isdir(DEST)
    n = 0
    for tree in TREES:
        for root, _dirs, files in os.walk(os.path.join(REF_SRC, tree)):
            for f in files:
                if not f.endswith(".py"):
                    continue
                src = os.path.join(root, f)
                dst = os.path.join(DEST, os.path.relpath(src, REF_SRC))
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                shutil.copyfile(src, dst)
                n += 1
    for f in FILES:
        dst = os.path.join(DEST, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF_SRC, f), dst)
        n += 1
    if verbose:
        print(f"[make_ref] copied {n} reference .py files into {DEST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if make_ref() else 1)
