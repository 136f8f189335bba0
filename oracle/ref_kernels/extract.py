#!/usr/bin/env python
"""Build step of oracle/_ref: cut the DEVICE code (templates, __device__ helpers, __global__
kernels) out of the reference's own .cu files, where they lie under the reference checkout, into
oracle/_ref/*.inc. Nothing of the reference is checked into this repository: oracle/_ref/ is
git-ignored, and only the line ranges below (with a token each range must start with, so that a
different reference revision fails loudly instead of cutting the wrong lines) are committed.

The host halves of those files (init_kernel, the DataTypeDispatch functors, forward_kernel /
backward_kernel) need the reference's generated headers (*.dtg.h) and cuDNN/cuBLAS handles and are
NOT extracted; oracle/ref_kernels/harness.cu restates their launch configurations, citing the
lines. Test infrastructure only."""
import sys
from pathlib import Path

SRC = "lib/kernels/src/cuda"
# name -> (file, [(first line, last line, token expected on the first line)])
RANGES = {
    "layer_norm": (f"{SRC}/ops/layer_norm_kernels.cu", [(22, 25, "#define C10_WARP_SIZE"), (30, 289, "template")]),
    "embedding": (f"{SRC}/embedding_kernels.cu", [(40, 342, "template")]),
    "loss": (f"{SRC}/loss_function_kernels.cu", [(21, 57, "__global__")]),
    "optimizer": (f"{SRC}/optimizer_kernel.cu", [(20, 41, "__global__ void sgd_update"), (110, 145, "__global__")]),
    "gather": (f"{SRC}/ops/gather_kernels.cu", [(25, 78, "template")]),
    "transpose": (f"{SRC}/ops/transpose_kernels.cu", [(23, 27, "struct TransposeStrides"), (45, 60, "__global__")]),
    "reverse": (f"{SRC}/ops/reverse_kernels.cu", [(24, 38, "__global__")]),
    "cast": (f"{SRC}/ops/cast_kernels.cu", [(24, 37, "template")]),
    "helper": (f"{SRC}/cuda_helper.cu", [(32, 63, "__global__ void scale_kernel"), (84, 90, "template"),
                                         (111, 172, "__global__ void gelu_forward_kernel")]),
    "element_unary": (f"{SRC}/ops/element_unary_kernels.cu", [(98, 245, "template")]),
    "metrics": (f"{SRC}/metrics_functions.cu", [(21, 137, "float const LOG_MIN_VALUE")]),
    "topk": (f"{SRC}/ops/topk_kernels.cu", [(24, 31, "enum class HeapType"), (38, 370, "template"),
                                            (409, 422, "template")]),
    "replicate": (f"{SRC}/ops/replicate_kernels.cu", [(24, 34, "template")]),
    "reduction": (f"{SRC}/ops/reduction_kernels.cu", [(24, 35, "template")]),
}


def main(ref_root: str, out_dir: str) -> int:
    ref, out = Path(ref_root), Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    for name, (rel, ranges) in RANGES.items():
        lines = (ref / rel).read_text().splitlines()
        chunks = [f"// cut from {rel} (build artefact, not tracked)"]
        for first, last, token in ranges:
            if token not in lines[first - 1]:
                print(f"extract: {rel}:{first} does not start with {token!r} — a different reference revision?",
                      file=sys.stderr)
                return 1
            chunks.append(f"#line {first} \"{rel}\"")
            chunks += lines[first - 1:last]
        (out / f"{name}.inc").write_text("\n".join(chunks) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1], sys.argv[2]))
