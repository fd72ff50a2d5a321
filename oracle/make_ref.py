"""Recipe for oracle/_ref: the reference's OWN files for the hot path, copied unmodified from where
they lie under /root/reference so that they travel to the GPU box (oracle/_ref/ is git-ignored, not
gpurun-ignored -- like a built .so; nothing under it is ever committed).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run by __graft_entry__.build() in the build
container; bench.py's CPU arms (`cpu_baseline`, `--impl reference`) then time these files through
oracle/reference_loader.py (`kind: "reference"`) and fall back to the numpy port only when the
directory is absent.  The three files are SURVEY.md 8(a): process/extract_rosbag_lidar.py (a1, a2),
common/birds_eye_view_generator.py (a3, a4), process/globals.py (a5).
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("SFPE_REFERENCE_SRC", "/root/reference")
FILES = ("modules/lidar/process/extract_rosbag_lidar.py", "modules/lidar/process/globals.py",
         "modules/lidar/common/birds_eye_view_generator.py")


def make(dst=os.path.join(HERE, "_ref")):
    """Returns the list of files copied ([] when the reference tree is not mounted)."""
    if not all(os.path.isfile(os.path.join(SRC, f)) for f in FILES):
        return []
    done = []
    for f in FILES:
        out = os.path.join(dst, f)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, f), out)
        done.append(out)
    return done


if __name__ == "__main__":
    got = make()
    print("oracle/_ref: %d files" % len(got) if got else "oracle/_ref: reference tree not mounted, nothing copied")
    sys.exit(0)
