"""Build libydorb.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import pathlib
import subprocess
import sys

HERE = pathlib.Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "libydorb.so"
OBJ = HERE / "_obj"
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
]


def sources():
    return sorted(CSRC.glob("*.cu"))


def needs_build():
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*")) + [HERE.parent / "include" / "ydorb.h"]
    return any(p.stat().st_mtime > t for p in deps)


def build(force=False, verbose=False):
    """One object per translation unit (no relocatable device code: no device call crosses a file), compiled in
    parallel and re-used while neither the source nor any header changed; then one link."""
    if not force and not needs_build():
        return LIB
    from concurrent.futures import ThreadPoolExecutor

    OBJ.mkdir(exist_ok=True)
    headers = [p for p in CSRC.glob("*") if p.suffix != ".cu"] + [HERE.parent / "include" / "ydorb.h"]
    hdr_t = max(p.stat().st_mtime for p in headers)

    def compile_one(src):
        obj = OBJ / (src.stem + ".o")
        if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_t):
            return obj
        cmd = ["nvcc", *NVCC_FLAGS, "-c", "-o", str(obj), str(src)]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
        return obj

    with ThreadPoolExecutor(max_workers=8) as pool:
        objs = list(pool.map(compile_one, sources()))
    cmd = ["nvcc", *NVCC_FLAGS, "-shared", "-o", str(LIB), *[str(o) for o in objs]]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print(LIB)
