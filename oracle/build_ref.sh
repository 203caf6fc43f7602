#!/usr/bin/env bash
# Builds "oracle A": the UNMODIFIED reference solver, compiled from the sources where they lie
# under /root/reference, into oracle/_ref/ (git-ignored; travels to the GPU box with gpurun).
#   libref_sim_f64.so  Real=double (the reference default, simulation.h:21-25)
#   libref_sim_f32.so  Real=float  (-DUSE_FLOATS via CUSTOM_SETTINGS, simulation.h:10-19)
#   libref_cg.so       the reference's CPU conjugate-gradient scratch program (cojugate_gradient.c),
#                      its main() renamed so the solver functions can be called for known answers
#   ref_cg_main        the same program as an executable (prints its own iteration counts)
# Test infrastructure only; no reference source is copied into this repository.
set -euo pipefail
REF=${REF_ROOT:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
mkdir -p "$OUT"
if [ ! -f "$REF/simulation.cu" ]; then
  echo "build_ref: $REF not present (GPU box) - using prebuilt files in $OUT" >&2
  exit 0
fi
NVCC=${NVCC:-nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
DEV_FLAGS="--extended-lambda --expt-relaxed-constexpr -DNDEBUG -std=c++17 -O3 -Xcompiler -fPIC -w"

build_sim () { # $1 = suffix, $2.. = extra defines
  local sfx=$1; shift
  if [ "$OUT/libref_sim_$sfx.so" -nt "$REF/simulation.cu" ] && [ "$OUT/libref_sim_$sfx.so" -nt "$HERE/ref_glue/ref_glue.cpp" ]; then return; fi
  # run from the reference directory: -I$REF would shadow <assert.h> with the reference's own assert.h
  if [ ! "$OUT/simulation_$sfx.o" -nt "$REF/simulation.cu" ]; then
    (cd "$REF" && $NVCC $ARCH $DEV_FLAGS "$@" -c simulation.cu -o "$OUT/simulation_$sfx.o")
  fi
  $NVCC $ARCH -std=c++17 -O2 -Xcompiler -fPIC -w "$@" -Xcompiler -iquote -Xcompiler "$REF" -I"$HERE/../include" \
      -c "$HERE/ref_glue/ref_glue.cpp" -o "$OUT/ref_glue_$sfx.o"
  $NVCC $ARCH -shared -o "$OUT/libref_sim_$sfx.so" "$OUT/simulation_$sfx.o" "$OUT/ref_glue_$sfx.o" -lcudart
  rm -f "$OUT/ref_glue_$sfx.o"
}

# The drop-in check: this repository's simulation_shim.cpp (the replacement of the reference's simulation.cu) behind the
# SAME glue and the reference's own simulation.h / log.h, linked against libpfsolver.so instead of the reference solver:
#   libshim_sim_f64.so / libshim_sim_f32.so  - the reference's structs and entry points at run time, our kernels below.
build_shim () { # $1 = suffix, $2.. = extra defines
  local sfx=$1; shift
  local shim="$HERE/../bachelors_b200/host/simulation_shim.cpp"
  if [ "$OUT/libshim_sim_$sfx.so" -nt "$shim" ] && [ "$OUT/libshim_sim_$sfx.so" -nt "$HERE/ref_glue/ref_glue.cpp" ] \
     && [ "$OUT/libshim_sim_$sfx.so" -nt "$HERE/../include/pf_solver.h" ]; then return; fi
  g++ -std=c++17 -O2 -fPIC -w "$@" -iquote "$REF" -I"$HERE/../include" -c "$shim" -o "$OUT/shim_$sfx.o"
  $NVCC $ARCH -std=c++17 -O2 -Xcompiler -fPIC -w "$@" -Xcompiler -iquote -Xcompiler "$REF" -I"$HERE/../include" \
      -c "$HERE/ref_glue/ref_glue.cpp" -o "$OUT/shim_glue_$sfx.o"
  g++ -shared -o "$OUT/libshim_sim_$sfx.so" "$OUT/shim_$sfx.o" "$OUT/shim_glue_$sfx.o" \
      -L"$HERE/../bachelors_b200" -lpfsolver -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,'$ORIGIN/../../bachelors_b200'
  rm -f "$OUT/shim_$sfx.o" "$OUT/shim_glue_$sfx.o"
}

# CPU CG program (fast)
if [ ! "$OUT/libref_cg.so" -nt "$REF/cojugate_gradient.c" ]; then
  g++ -std=c++17 -O2 -w -fPIC -shared -Dmain=ref_cg_main -x c++ "$REF/cojugate_gradient.c" -o "$OUT/libref_cg.so"
  g++ -std=c++17 -O2 -w -x c++ "$REF/cojugate_gradient.c" -o "$OUT/ref_cg_main"
fi

build_sim f64 &
build_sim f32 -DCUSTOM_SETTINGS -DUSE_FLOATS -DCOMPILE_SIMULATION &
if [ -f "$HERE/../bachelors_b200/libpfsolver.so" ]; then
  build_shim f64 &
  build_shim f32 -DCUSTOM_SETTINGS -DUSE_FLOATS -DCOMPILE_SIMULATION &
fi
wait
ls -la "$OUT"
