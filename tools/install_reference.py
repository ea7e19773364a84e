"""Copies the UNMODIFIED reference sources the CPU baseline runs (src/grpc_node.py, src/run_grpc_inference.py,
src/proto/*) from /root/reference into the git-ignored baseline/_ref/, so that they travel to the GPU box with
the repo snapshot (the box has no /root/reference).  Called by __graft_entry__.build() in the build container;
nothing in the product or in the GPU arm of bench.py imports the copy -- only `bench.py --impl reference`
and tools/reference_chain.py do.   python tools/install_reference.py
"""
import os
import shutil
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/src"
DST = os.path.join(REPO, "baseline", "_ref", "src")
FILES = ["grpc_node.py", "run_grpc_inference.py", "proto/__init__.py", "proto/dist_nn.proto",
         "proto/dist_nn_pb2.py", "proto/dist_nn_pb2_grpc.py"]


def install() -> bool:
    """True when baseline/_ref holds the reference files (copied now or earlier)."""
    if not os.path.isdir(SRC):
        return os.path.exists(os.path.join(DST, "grpc_node.py"))
    for f in FILES:
        dst = os.path.join(DST, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, f), dst)
    return True


if __name__ == "__main__":
    ok = install()
    print(DST if ok else "reference sources not available")
    sys.exit(0 if ok else 1)
