#!/bin/bash
./gpu_ab.sh
OUT=${OUT:-prof_k1e} ./gpu_prof.sh
