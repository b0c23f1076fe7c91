#!/bin/bash
# TEST INFRASTRUCTURE ONLY.  Builds the UNMODIFIED reference (its .cu sources compiled where they lie under /root/reference,
# nothing copied) plus our small driver oracle/ref_harness.cu into oracle/_ref/ref_harness: the reference's own CUDA
# implementation of the path, cross-compiled for sm_100 here and executed only on the GPU box by tests/ and bench.py as a
# checker / baseline.  This is a short hand-written recipe, not the reference's CMake build.  Outputs go to oracle/_ref/ only
# (git-ignored, shipped to the GPU box by gpurun).
#
#   -D__CUDA_GL_INTEROP_H__   internal/utils/render_buffer.cu includes <cuda_gl_interop.h> but uses no GL symbol; defining the
#                             header's include guard skips it (there is no graphics stack on these machines and none is needed)
#   -include cstdint          tcnn headers of this vintage rely on transitive <cstdint>
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
[ -d "$REF" ] || { echo "build_ref: $REF not present (GPU box uses the prebuilt oracle/_ref)"; exit 0; }
mkdir -p "$OUT/obj"
TCNN=$REF/dependencies/tiny-cuda-nn
NVCC=${NVCC:-nvcc}
FLAGS=(-gencode arch=compute_100,code=sm_100 -std=c++14 -O3 --extended-lambda --expt-relaxed-constexpr -rdc=true -w
       -DTCNN_AMPERE -DTCNN_SHAMPOO -DNGP_VERSION=\"1.0\" -D__CUDA_GL_INTEROP_H__ -include cstdint
       -Xcompiler=-mf16c,-fno-strict-aliasing,-fPIC,-fopenmp
       -I"$REF" -I"$REF/dependencies" -I"$REF/dependencies/eigen" -I"$REF/dependencies/filesystem" -I"$REF/dependencies/tinylogger"
       -I"$TCNN/include" -I"$TCNN/dependencies" -I"$TCNN/dependencies/cutlass/include" -I"$TCNN/dependencies/cutlass/tools/util/include")
SRCS=("$TCNN"/src/{common,common_device,cutlass_mlp,cutlass_resnet,encoding,fully_fused_mlp,loss,network,object,optimizer,reduce_sum}.cu
      "$REF"/internal/dataset/nerf_synthetic.cu "$REF"/internal/sampler/occupancy_sampler.cu "$REF"/internal/render/ray_marcher.cu
      "$REF"/internal/utils/{common_device,loss,render_buffer}.cu "$REF"/internal/testbed.cu "$HERE/ref_harness.cu")
JOBS=${JOBS:-8}
compile_one() {
	local src="$1" tag obj
	tag=$(echo "$src" | sed -e "s#$REF/##" -e "s#$HERE/##" -e 's#[/.]#_#g')
	obj="$OUT/obj/$tag.o"
	if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ]; then
		echo "  nvcc $src"
		"$NVCC" "${FLAGS[@]}" -dc "$src" -o "$obj" 2> "$obj.log" || { echo "FAILED $src (see $obj.log)"; tail -20 "$obj.log"; return 1; }
	fi
}
pids=()
for s in "${SRCS[@]}"; do
	[ -f "$s" ] || continue
	compile_one "$s" &
	pids+=($!)
	while [ "$(jobs -rp | wc -l)" -ge "$JOBS" ]; do sleep 0.5; done
done
fail=0
for p in "${pids[@]}"; do wait "$p" || fail=1; done
[ $fail = 0 ] || { echo "build_ref: compile failed"; exit 1; }
[ -f "$HERE/ref_harness.cu" ] || { echo "build_ref: objects built, no harness yet"; exit 0; }
"$NVCC" -gencode arch=compute_100,code=sm_100 -Xcompiler=-fopenmp "$OUT"/obj/*.o -o "$OUT/ref_harness" -lcublas -lcuda
echo "built $OUT/ref_harness"
