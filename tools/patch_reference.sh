#!/bin/bash
# tools/patch_reference.sh -- ONE harness binary that dispatches kernel numbers 0..21.
#
# Copies the reference tree to a scratch directory OUTSIDE this repo (reference sources are never copied into the repo),
# applies the 4-file patch of INTEGRATION.md to the copy (kernels.h, utils.c, test.c, Makefile), and builds it with the
# reference's own flags (reference Makefile:3) against libdgemm_b200.so, with cblas_dgemm (kernel 0 and the harness's own
# verification call, test.c:38) served by harness/mkl_b200/mkl.h because MKL does not exist on this box.  Only the binary
# comes back: oracle/_ref/dgemm_x86_patched (git-ignored, travels to the GPU box like the other oracle/_ref files).
#
#   tools/patch_reference.sh [REF_ROOT=/root/reference] [SCRATCH=/tmp/dgemm_ref_patched]
#   oracle/_ref/dgemm_x86_patched 20        # the reference's harness, kernel 20 = mydgemm_gpu_v20
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF_ROOT="${1:-/root/reference}"
SCRATCH="${2:-/tmp/dgemm_ref_patched}"
PKG="$REPO/optimizing-dgemm-on-intel-cpus-with-avx512f_b200"
[ -f "$REF_ROOT/kernels.h" ] || { echo "patch_reference: $REF_ROOT has no kernels.h" >&2; exit 2; }
[ -f "$PKG/libdgemm_b200.so" ] || { echo "patch_reference: build the library first (make lib)" >&2; exit 2; }
rm -rf "$SCRATCH"; mkdir -p "$SCRATCH"
cp "$REF_ROOT"/test.c "$REF_ROOT"/utils.c "$REF_ROOT"/utils.h "$REF_ROOT"/kernels.h "$REF_ROOT"/Makefile "$SCRATCH"/
cp -r "$REF_ROOT"/include "$SCRATCH"/include
cd "$SCRATCH"
# --- kernels.h: one more header behind the nineteen (reference kernels.h:1-19)
printf '\n#include "kernel20.h"   /* mydgemm_gpu_v20 / v21 + test_mydgemm_v20 / v21 (this repo, include/kernel20.h) */\n' >> kernels.h
# --- utils.c: two more cases in the switch of test_kernel (reference utils.c:167-191)
grep -q 'case 19: test_mydgemm_v19' utils.c
sed -i 's|^\(\s*\)case 19: test_mydgemm_v19(m,n,k,alpha,A,B,beta,C); break;|&\n\1case 20: test_mydgemm_v20(m,n,k,alpha,A,B,beta,C); break;\n\1case 21: test_mydgemm_v21(m,n,k,alpha,A,B,beta,C); break;\n\1case 22: case 23: case 24: case 25: case 26: case 27: test_mydgemm_ladder(kernel_num,m,n,k,alpha,A,B,beta,C); break;|' utils.c
grep -q 'case 21: test_mydgemm_v21' utils.c
# --- test.c: accepted range 0..21 (reference test.c:9, 16-17); fail loudly when a GPU kernel reports an error; optional
#     cap of the size sweep for automated runs (DGEMM_X86_MAX_SIZE)
sed -i 's|range 0 - 19|range 0 - 27|; s|kernel_num>19|kernel_num>27|; s|valid kernel number (0-19)|valid kernel number (0-27)|' test.c
sed -i 's|^\(\s*\)m=n=k=SIZE\[i_count\];|\1m=n=k=SIZE[i_count];\n\1if (getenv("DGEMM_X86_MAX_SIZE") \&\& m>atoi(getenv("DGEMM_X86_MAX_SIZE"))) break;|' test.c
sed -i 's|^\(\s*\)t0=get_sec();|\1if (kernel_num>=20 \&\& dgemm_b200_last_status()!=0){ printf("GPU kernel failed: %s\\n",dgemm_b200_last_error()); exit(-6); }\n&|' test.c
grep -q 'kernel_num>27' test.c && grep -q 'dgemm_b200_last_status' test.c
# --- Makefile: include path + library (reference Makefile:3-6); MKL's include / link lines give way to the stand-in
sed -i "s|^LDFLAGS.*|LDFLAGS\t\t= -L$PKG -Wl,--no-as-needed -ldgemm_b200 -Wl,-rpath,'\$\$ORIGIN/../../optimizing-dgemm-on-intel-cpus-with-avx512f_b200' -lpthread -lm -ldl|; s|^INCFLAGS.*|INCFLAGS\t= -I$PKG/harness/mkl_b200 -I$REPO/include|; s|^CC\t\t\t= gcc|CC\t\t\t= /usr/bin/gcc|" Makefile
make build > build.log 2>&1 || { cat build.log >&2; exit 3; }
mkdir -p "$REPO/oracle/_ref"
cp dgemm_x86 "$REPO/oracle/_ref/dgemm_x86_patched"
echo "patch_reference: built $REPO/oracle/_ref/dgemm_x86_patched (kernel numbers 0..27); patched copy in $SCRATCH"
# --- the same patched tree once more, with cblas_dgemm served by the CHECKER (oracle/mkl_shim.c -> the kernel18 restatement
#     in liboracle.so) instead of by the B200 library: in THIS binary the harness's own 1e-2 verification of kernels 20 / 21
#     (reference test.c:36-43) compares the GPU with an independent CPU result, not the GPU with itself
if [ -f "$REPO/oracle/liboracle.so" ]; then
  cp "$REPO/oracle/mkl_shim.c" "$REPO/oracle/mkl.h" .
  rm -f dgemm_x86
  make build INCFLAGS="-I. -I$REPO/include" \
       LDFLAGS="-L$PKG -L$REPO/oracle -Wl,--no-as-needed -ldgemm_b200 -loracle -Wl,-rpath,'\$\$ORIGIN/../../optimizing-dgemm-on-intel-cpus-with-avx512f_b200' -Wl,-rpath,'\$\$ORIGIN/..' -lpthread -lm -ldl" \
       > build_oracle.log 2>&1 || { cat build_oracle.log >&2; exit 3; }
  cp dgemm_x86 "$REPO/oracle/_ref/dgemm_x86_patched_oracle"
  rm -f mkl_shim.c mkl.h
  echo "patch_reference: built $REPO/oracle/_ref/dgemm_x86_patched_oracle (cblas_dgemm = the plain-C oracle)"
fi
diff -u "$REF_ROOT/test.c" test.c > "$SCRATCH/test.c.diff" || true
diff -u "$REF_ROOT/utils.c" utils.c > "$SCRATCH/utils.c.diff" || true
