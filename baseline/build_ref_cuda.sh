#!/usr/bin/env bash
# BASELINE B1: the reference's own CUDA walker (elaina-exec), built out of tree from the sources where they
# lie under /root/reference.  TEST / BASELINE INFRASTRUCTURE ONLY -- nothing under elaina_b200/ uses it.
#
#   * nothing is copied into the repository: objects go to $WORK (/tmp), the binary to baseline/_ref/
#     (git-ignored, travels to the GPU box with the gpurun snapshot);
#   * the reference's cmake is not run: this script is the whole recipe (soac generator -> 4 SoA headers,
#     tiny-cuda-nn static objects, fmt, the elaina core/util/integrator translation units, one link);
#   * OpenVDB (TBB/Boost absent) is not needed: the path only uses the header-only NanoVDB tree;
#   * Catch2 tests are not built (ext/Catch2 is empty in the snapshot);
#   * arch: sm_100 SASS (tcnn's cmake caps the list at 90 only because of its CUDA-version table;
#     TCNN_MIN_GPU_ARCH=100 compiles, its kernels are plain wmma);
#   * TWO instrumentation patches, applied to a scratch copy of the two integrator.cu files: (i) with B1_DUMP_RAW=<path>
#     in the environment the solution film {r,g,b,a,weight} x W*H is written as raw fp32 after solve() (the reference's
#     own .exr export is half precision); (ii) a managed counter
#     summing the evaluation-point queue size per depth inside the reference's own per-depth reset kernel
#     (walk-steps, the unit of BASELINE.json's metric), printed after solve().
#     No extra launch, no extra synchronisation.
# Flags follow CMakeLists.txt:64-70 (--use_fast_math --expt-relaxed-constexpr --extended-lambda, -O3 Release).
set -u
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
WORK=${WORK:-/tmp/elaina_b1}
OUT=$HERE/_ref
JOBS=${JOBS:-8}
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH=${ARCH:-"-gencode arch=compute_100,code=sm_100"}
MINARCH=${MINARCH:-100}
if [ ! -d "$REF/integrator" ]; then echo "reference tree absent: keeping prebuilt baseline/_ref (if any)"; exit 0; fi
mkdir -p "$WORK/obj" "$WORK/gen/integrator/uniform" "$WORK/gen/integrator/guided" "$WORK/patched/uniform" "$WORK/patched/guided" "$OUT"
LOG=$WORK/build.log; : > "$LOG"
say() { echo "$@" | tee -a "$LOG"; }
run() { say "+ $*"; "$@" >>"$LOG" 2>&1; local rc=$?; [ $rc -ne 0 ] && say "FAILED($rc): $*"; return $rc; }

# 1. SoA generator and its four generated headers (CMakeLists.txt:186-236)
run g++ -std=c++17 -O1 -o "$WORK/soac" "$REF/util/soac.cpp" || exit 1
for p in uniform/uniform2d uniform/uniform3d guided/guided2d guided/guided3d; do
  "$WORK/soac" "$REF/integrator/${p}_workitem.soa" > "$WORK/gen/integrator/${p}_workitem_soa.h" || { say "soac failed on $p"; exit 1; }
done

# 2. walk-step counter patch on scratch copies
for k in uniform guided; do
  python3 - "$REF/integrator/$k/integrator.cu" "$WORK/patched/$k/integrator.cu" <<'EOF' || exit 1
import re, sys
src = open(sys.argv[1]).read()
# (a) the counter, after the namespace opens
src = src.replace("ELAINA_NAMESPACE_BEGIN", "ELAINA_NAMESPACE_BEGIN\n__device__ __managed__ unsigned long long g_b1_walk_steps = 0ull;\n", 1)
# (b) count inside the per-depth reset kernel of solveImpl
needle = "ctx.nextEvaluationPointQueue(depth)->reset();"
assert src.count(needle) == 1, src.count(needle)
src = src.replace(needle, "g_b1_walk_steps += (unsigned long long)ctx.currentEvaluationPointQueue(depth)->size();\n                    " + needle)
# (c) report after the final sync of solveImpl
needle = "ELAINA_DESTROY_PROGRESS_BAR();"
i = src.rfind(needle, 0, src.find("resetNetworkImpl") if "resetNetworkImpl" in src else len(src))
if i < 0: i = src.rfind(needle)
dump = ('if (const char *b1_raw = getenv("B1_DUMP_RAW")) { auto *b1_film = ctx.get_solutionChannel(); '
        'const size_t b1_n = (size_t)b1_film->size()[0] * (size_t)b1_film->size()[1]; '
        'std::vector<Film::WeightedPixel> b1_h(b1_n); '
        'cudaMemcpy(b1_h.data(), b1_film->data(), b1_n * sizeof(Film::WeightedPixel), cudaMemcpyDefault); '
        'if (FILE *b1_f = fopen(b1_raw, "wb")) { for (size_t b1_i = 0; b1_i < b1_n; ++b1_i) { '
        'float b1_v[5] = {b1_h[b1_i].pixel[0], b1_h[b1_i].pixel[1], b1_h[b1_i].pixel[2], b1_h[b1_i].pixel[3], b1_h[b1_i].weight}; '
        'fwrite(b1_v, sizeof(float), 5, b1_f); } fclose(b1_f); } }\n        ')
src = src[:i] + dump + 'printf("B1_WALK_STEPS %llu\\n", g_b1_walk_steps); fflush(stdout);\n        ' + src[i:]
open(sys.argv[2], "w").write(src)
EOF
done

TCNN=$REF/ext/tcnn
INC_TCNN="-I$TCNN/include -I$TCNN/dependencies -I$TCNN/dependencies/cutlass/include -I$TCNN/dependencies/cutlass/tools/util/include -I$TCNN/dependencies/fmt/include"
DEFS="-DTCNN_MIN_GPU_ARCH=$MINARCH -DTCNN_SHAMPOO -DFCPW_USE_GPU -DNDEBUG"
CUFLAGS="$ARCH -std=c++17 -O3 --use_fast_math --expt-relaxed-constexpr --extended-lambda -lineinfo -w \
 -Xcompiler=-mf16c -Xcompiler=-fno-strict-aliasing -Xcudafe=--diag_suppress=unrecognized_gcc_pragma $DEFS"
INC_ELAINA="-I$REF -I$WORK/gen -I$REF/ext/image -I$REF/ext/eigen -I$REF/ext/json/single_include $INC_TCNN \
 -I$REF/core/math/include -I$REF/ext/indicators/single_include/indicators -I$REF/ext/lbvh/include \
 -I$REF/ext/magic_enum/include -I$REF/ext/openvdb/nanovdb"

# 3. job list: "<object> <command...>"
JOBFILE=$WORK/jobs.txt; : > "$JOBFILE"
for f in common common_device cpp_api cutlass_mlp encoding loss network object optimizer reduce_sum fully_fused_mlp; do
  echo "$NVCC $CUFLAGS $INC_TCNN -c $TCNN/src/$f.cu -o $WORK/obj/tcnn_$f.o" >> "$JOBFILE"
done
for f in format os; do
  echo "g++ -std=c++17 -O2 -I$TCNN/dependencies/fmt/include -c $TCNN/dependencies/fmt/src/$f.cc -o $WORK/obj/fmt_$f.o" >> "$JOBFILE"
done
for f in core/file.cpp core/interop.cpp core/logger.cpp core/device/context.cpp util/image.cpp main.cpp; do
  o=$(echo "$f" | tr '/.' '__')
  echo "$NVCC $CUFLAGS $INC_ELAINA -x cu -c $REF/$f -o $WORK/obj/$o.o" >> "$JOBFILE"
done
for f in core/problem.cu core/texture.cu exec.cu; do
  o=$(echo "$f" | tr '/.' '__')
  echo "$NVCC $CUFLAGS $INC_ELAINA -c $REF/$f -o $WORK/obj/$o.o" >> "$JOBFILE"
done
for k in uniform guided; do
  echo "$NVCC $CUFLAGS -I$REF/integrator/$k $INC_ELAINA -c $WORK/patched/$k/integrator.cu -o $WORK/obj/integrator_$k.o" >> "$JOBFILE"
done
say "compiling $(wc -l < "$JOBFILE") translation units with $JOBS jobs (log: $LOG)"
export LOG
# shellcheck disable=SC2016
xargs -P "$JOBS" -I{} bash -c 'o=$(echo "{}" | sed "s/.* -o //"); if [ -s "$o" ] && [ -z "${FORCE:-}" ]; then exit 0; fi; t0=$(date +%s); out=$({} 2>&1); rc=$?; echo "[$(( $(date +%s) - t0 )) s, rc=$rc] {}" >> "$LOG"; [ $rc -ne 0 ] && echo "$out" | head -60 >> "$LOG"; exit 0' < "$JOBFILE"
missing=0
while read -r line; do o=$(echo "$line" | sed "s/.* -o //"); [ -s "$o" ] || { say "MISSING OBJECT: $o"; missing=1; }; done < "$JOBFILE"
[ $missing -ne 0 ] && { say "B1 build FAILED (see $LOG)"; exit 1; }

# 4. link (libcuda comes from the driver on the GPU box; the toolkit stub satisfies the linker here)
run $NVCC $ARCH -o "$OUT/elaina-exec-ref" "$WORK"/obj/*.o -L/usr/local/cuda/lib64 -L/usr/local/cuda/lib64/stubs -lcuda -lcublas \
  -Xlinker -rpath=/usr/local/cuda/lib64 || { say "B1 link FAILED"; exit 1; }
say "B1 build OK: $OUT/elaina-exec-ref ($(stat -c %s "$OUT/elaina-exec-ref") bytes)"
