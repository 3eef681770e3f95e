"""Compiles the reference's own iou3d_cpu.cpp (from /root/reference, where it lies) plus a small C-ABI shim
into oracle/_ref/libiou3d_ref.so. Test infrastructure; run by `make -C oracle` in the build container.
The .so is git-ignored and travels to the GPU box with the snapshot (same image, same libtorch)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("TCB_REFERENCE_ROOT", "/root/reference")
SRC = os.path.join(REF, "utils/export_models/OpenPCDet/pcdet/ops/iou3d_nms/src")


def main():
    import torch
    from torch.utils import cpp_extension
    out = os.path.join(HERE, "_ref")
    os.makedirs(out, exist_ok=True)
    import sysconfig
    inc = cpp_extension.include_paths() + [SRC, "/usr/local/cuda/include", sysconfig.get_paths()["include"]]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-w", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    cmd += [f"-I{i}" for i in inc]
    cmd += [os.path.join(SRC, "iou3d_cpu.cpp"), os.path.join(HERE, "ref_iou3d_bind.cpp")]
    cmd += ["-o", os.path.join(out, "libiou3d_ref.so"), f"-L{libdir}", "-ltorch", "-ltorch_cpu", "-lc10", f"-Wl,-rpath,{libdir}"]
    print(" ".join(cmd))
    subprocess.check_call(cmd)


if __name__ == "__main__":
    if not os.path.isfile(os.path.join(SRC, "iou3d_cpu.cpp")):
        print("reference not mounted; nothing to build")
        sys.exit(0)
    main()
