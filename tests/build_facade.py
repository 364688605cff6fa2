"""Builds tests/_build/libfacade_test.so: include/spm_cuda_processor.h against the unmodified
reference header, linked with oracle/_ref/libspm_ref.so and the product library.  Only possible
where /root/reference is mounted (the build container); the GPU box uses the prebuilt file."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("SPM_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "_build", "libfacade_test.so")
SRC = os.path.join(ROOT, "tests", "facade_driver.cc")
HDRS = [os.path.join(ROOT, "include", "spm_cuda_processor.h"), os.path.join(ROOT, "include", "spm_cuda.h")]


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "src")):
        return OUT if os.path.exists(OUT) else None
    ref_lib = os.path.join(ROOT, "oracle", "_ref")
    cuda_lib = os.path.join(ROOT, "sentencepiece-comments_b200", "lib")
    deps = [SRC] + HDRS + [os.path.join(ref_lib, "libspm_ref.so")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    subprocess.check_call([
        "g++", "-std=c++11", "-O1", "-shared", "-fPIC", "-Wall", "-Wno-deprecated-declarations",
        "-I" + os.path.join(REF, "src"), "-I" + os.path.join(REF, "src", "builtin_pb"),
        "-I" + os.path.join(REF, "third_party", "protobuf-lite"), "-I" + REF, "-I" + os.path.join(ROOT, "include"),
        SRC, "-o", OUT, "-L" + ref_lib, "-l:libspm_ref.so", "-L" + cuda_lib, "-l:libspm_cuda.so",
        "-Wl,-rpath,$ORIGIN/../../oracle/_ref", "-Wl,-rpath,$ORIGIN/../../sentencepiece-comments_b200/lib"])
    return OUT


if __name__ == "__main__":
    print(build(force=True))
