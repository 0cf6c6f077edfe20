#!/bin/bash
# (line ranges follow ray_tracer_b200/csrc/pt_traverse.h as of the final build of round 2)
# Groups a sass_by_line.py listing (all lines: pass a large `top`) by code region of the traversal kernels:
#   bash profiles/tools/by_region.sh <wavefront.sass> <kernel substring> <src.csv>
python "$(dirname "$0")/sass_by_line.py" "$1" "$2" "$3" 1000 2>&1 | awk 'NR>2 {split($1,a,":"); f=a[1]; l=a[2]+0; r="other";
 if (f=="pt_traverse.h") { if (l>=259 && l<=352) r="ref_reachable"; else if (l>=463 && l<=541) r="triangle step glue"; else if (l>=53 && l<=220) r="node test"; else if (l>=221 && l<=258) r="object_ray / load_prim"; else if (l>=353 && l<=462) r="stack, pop, node step glue"; else r="pt_traverse.h other" }
 else if (f=="wavefront.cu") r="kernel loop: votes, refill (wavefront.cu)"; else if (f=="pt_math.h") r="pt_math.h at " ($4<5?"< 5 lanes":($4<14?"5-14 lanes (triangle test)":"> 14 lanes (node test FMA)")); else if (f=="pt_shade.h") r="pt_shade.h (tri_test, spheres)"; else if (f=="pt_wavefront.h") r="pt_wavefront.h (spheres_closest, ray load)"; else r=f;
 W[r]+=$2; S[r]+=$3; T[r]+=$2*$4} END {for (r in W) printf "%5.1f%% of warp instructions  %5.1f%% of stall samples  %4.1f lanes  %s\n", W[r], S[r], T[r]/W[r], r}' | sort -n -r
