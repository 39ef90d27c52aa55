"""Builds build/emu/libdrm_emu.so: embed_kernel.cu and l2_kernel.cu compiled as C++ for the host (tools/cpu_emu/README.md)."""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "deepreadmapper_b200", "csrc")
OUT = os.path.join(ROOT, "build", "emu")
EMULATED = ("embed_kernel.cu", "l2_kernel.cu", "align_kernel.cu", "merge_kernel.cu", "input_kernels.cu", "index_image.cu", "search_kernel.cu",
            "search_exact.cu")
LAUNCH = re.compile(r"([A-Za-z_]\w*(?:<[^<>;]*>)?)\s*<<<(.*?)>>>\s*\((.*?)\);", re.S)


def transform(src: str) -> str:
    def repl(m):
        name, cfg, args = m.group(1), m.group(2), m.group(3)
        parts = [p.strip() for p in cfg.split(",")]
        smem = parts[2] if len(parts) > 2 else "0"
        return f"emu::launch(dim3({parts[0]}), dim3({parts[1]}), (size_t)({smem}), [&] {{ {name}({args}); }});"
    out, n = LAUNCH.subn(repl, src)
    if n == 0:
        raise SystemExit("no kernel launch found")
    # dynamic shared memory: `extern __shared__ T name[];` becomes a pointer to the block's buffer
    out = re.sub(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?([\w ]+?)\s+(\w+)\[\];", r"\1 *\2 = static_cast<\1 *>(emu::dyn_smem());", out)
    return out


def build(sanitize=True) -> str:
    os.makedirs(OUT, exist_ok=True)
    srcs = []
    for f in EMULATED:
        text = transform(open(os.path.join(CSRC, f)).read())
        dst = os.path.join(OUT, f.replace(".cu", "_emu.cpp"))
        with open(dst, "w") as fh:
            fh.write(text)
        srcs.append(dst)
    # the C-ABI layer itself (no kernel launches of its own) and the plain C++ files it needs
    api = os.path.join(OUT, "api_emu.cpp")
    with open(api, "w") as fh:
        fh.write(open(os.path.join(CSRC, "api.cu")).read())
    srcs += [api] + [os.path.join(CSRC, f) for f in ("ir_model.cpp", "l2_plan.cpp", "index_io.cpp", "builder.cpp")]
    srcs += [os.path.join(HERE, "emu_driver.cpp"), os.path.join(HERE, "emu_stubs.cpp")]
    lib = os.path.join(OUT, "libdrm_emu_asan.so" if sanitize else "libdrm_emu.so")
    newest = max(os.path.getmtime(p) for p in srcs + [os.path.join(HERE, "cuda_runtime.h"), os.path.join(CSRC, "common.h")])
    if os.path.exists(lib) and os.path.getmtime(lib) >= newest:
        return lib
    cmd = ["/usr/bin/g++", "-std=c++20", "-O1", "-g", "-fPIC", "-shared", "-pthread", "-fopenmp", "-DDRM_EMU", "-Wl,--no-undefined", "-I", HERE, "-I", CSRC,
           "-o", lib] + srcs
    if sanitize:
        cmd[3:3] = ["-fsanitize=address", "-fno-omit-frame-pointer"]
    subprocess.check_call(cmd)
    return lib


if __name__ == "__main__":
    print(build(sanitize="--no-asan" not in sys.argv))
