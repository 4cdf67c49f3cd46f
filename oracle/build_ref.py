"""TEST / BASELINE INFRASTRUCTURE ONLY — never imported by the product path (cp2_b200/).

Build recipe for `oracle/_ref/`: byte-compiles the reference's own modules of the CP2
pre-training path FROM THE SOURCES WHERE THEY LIE under /root/reference into sourceless
byte-code files `*.pycode` (a `.pyc` under another extension: the GPU-box snapshot drops `*.pyc`;
the Python analogue of compiling a C reference into `oracle/_ref/*.so`).
No reference source is copied into the repo; `oracle/_ref/` is git-ignored and is not
gpurun-ignored, so the compiled modules travel to the GPU box, where
`bench.py --impl reference` (CPU arm) and the `gpu_eager_baseline` leg import the
UNMODIFIED reference `builder.MODEL` through `oracle/ref_harness.py`.

    python oracle/build_ref.py            # also run by __graft_entry__.build() when /root/reference exists

Modules compiled (everything `import builder` pulls in from the reference tree, plus the mirror variant):
    builder.py, networks/__init__.py, networks/segment_network.py, networks/mirror_network.py,
    datasets/__init__.py, datasets/pretrain_dataset.py, tools/correlation_mapping.py
"""
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
MODULES = ["builder.py", "networks/__init__.py", "networks/segment_network.py", "networks/mirror_network.py",
           "datasets/__init__.py", "datasets/pretrain_dataset.py", "tools/correlation_mapping.py"]


def build(reference_root="/root/reference", out=OUT, quiet=False):
    """-> list of compiled files ([] when the reference tree is absent, e.g. on the GPU box)."""
    if not os.path.isfile(os.path.join(reference_root, "builder.py")):
        return []
    done = []
    for rel in MODULES:
        src = os.path.join(reference_root, rel)
        if not os.path.isfile(src):
            continue
        dst = os.path.join(out, rel[:-3] + ".pycode")
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        # unchecked-hash pyc: valid without the source file next to it (sourceless import)
        py_compile.compile(src, cfile=dst, dfile=rel, doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
        done.append(dst)
    with open(os.path.join(out, "README.txt"), "w") as f:
        f.write("Byte-compiled (sourceless) modules of the reference, built by oracle/build_ref.py from "
                f"{reference_root} with Python {sys.version.split()[0]}.\nBuild artefact: git-ignored, not source.\n")
    if not quiet:
        print(f"oracle/_ref: compiled {len(done)} reference modules from {reference_root}")
    return done


if __name__ == "__main__":
    build(*(sys.argv[1:2]))
