"""Tuning builds of libctrhot.so with different ring depths of the whole-model kernel (csrc/fused.cu).

    python profiles/build_variants.py NR=3,NB=3 NR=2,NB=4 ...   -> recommender_system_b200/build/variants/libctrhot_<tag>.so
    CTRHOT_LIB=<that file> python bench.py ...

Not part of the product: the shipped library is the default build (`python -m recommender_system_b200.build`).
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200 import build as B  # noqa: E402

OUT = os.path.join(B.OBJ, "variants")


def main(specs):
    os.makedirs(OUT, exist_ok=True)
    B.build_lib(verbose=False)
    src_name = "fused.cu"
    if specs and specs[0].startswith("src="):
        src_name, specs = specs[0][4:], specs[1:]
    others = [os.path.join(B.OBJ, s[:-3] + ".o") for s in B._sources() if s != src_name]
    for spec in specs:
        tag = spec.replace("=", "").replace(",", "_")
        defs = [("-D" + kv[2:]) if kv.startswith("D:") else (("-DFU_TRACE" if src_name == "fused.cu" else "-DDT_TRACE") if kv == "TRACE=1" else "-DDT_FIXED_ORDER" if kv == "FIXED=1" else "-DDT_OLD_WSUM" if kv == "OLDW=1" else "-DDT_LATE_ARRIVE" if kv == "LATE=1" else "-DDT_OLD_STAGER" if kv == "OLDS=1"
                 else "-DFU_%s_=%s" % tuple(kv.split("="))) for kv in spec.split(",")]
        tag = tag.replace("D:", "")
        obj = os.path.join(OUT, "%s_%s.o" % (src_name[:-3], tag))
        tag = src_name[:-3] + "_" + tag if src_name != "fused.cu" else tag
        lib = os.path.join(OUT, "libctrhot_%s.so" % tag)
        subprocess.check_call([B.NVCC] + B.FLAGS + defs + ["-c", os.path.join(B.CSRC, src_name), "-o", obj])
        subprocess.check_call([B.NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib, obj] + others)
        print(lib)


if __name__ == "__main__":
    main(sys.argv[1:])
