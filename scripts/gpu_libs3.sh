#!/bin/bash
bash scripts/gpu_libs.sh "$1"
unset LBCHEM_LIB LBCHEM_VARIANT
bash scripts/gpu_ncu2.sh g2 "--workload gri_uniform --solver seulex --cells 100000"
