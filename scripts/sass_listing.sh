#!/bin/bash
# SASS of the hot kernels (committed evidence, north_star): run after building the library.
#   bash scripts/sass_listing.sh  ->  profiles/sass_heat_step_vec.txt, profiles/sass_heat_step_tb.txt
set -e
LIB=parallel_heat_b200/libheat_b200.so
mkdir -p profiles
dump() { # $1 = substring of the mangled name, $2 = output
  cuobjdump -sass $LIB 2>/dev/null | awk -v pat="$1" '/Function : /{f = index($0, pat) > 0} f' > "$2"
  echo "$2: $(grep -cE '^\s+/\*[0-9a-f]{4}\*/' "$2") instructions"
  echo "# instruction mix" >> "$2"
  grep -E '^\s+/\*[0-9a-f]{4}\*/' "$2" | awk '{print $2}' | sed 's/\..*//;s/;//' | sort | uniq -c | sort -rn | head -25 | sed 's/^/# /' >> "$2"
}
dump "heat_step_vecILi0ELb0ELi128ELi4ELb0ELb0" profiles/sass_heat_step_vec_fma.txt
dump "heat_step_vecILi0ELb1ELi128ELi4ELb0ELb0" profiles/sass_heat_step_vec_fma_residual.txt
dump "heat_step_tbILi0ELi4ELi1ELb0" profiles/sass_heat_step_tb_fma_T4.txt
dump "heat_step_bulkILi0ELb0ELi8ELi128ELb0ELb0" profiles/sass_heat_step_bulk_fma.txt
dump "heat_step_vecILi2ELb0ELi128ELi4ELb0ELb0" profiles/sass_heat_step_vec_f64mix.txt
dump "heat_step_f64ILb1ELb0ELb0ELb0" profiles/sass_heat_step_f64_fma.txt
grep -cE "UTMALDG|UBLKCP|LDG.E.128|STG.E.128|SHFL|REDUX" profiles/sass_heat_step_vec_fma_residual.txt || true
