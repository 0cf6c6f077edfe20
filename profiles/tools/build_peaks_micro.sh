#!/bin/bash
# builds profiles/tools/peaks_micro (git-ignored binary) for sm_100a
cd "$(dirname "$0")" && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peaks_micro peaks_micro.cu
