#!/bin/sh
# SURVEY 7.1 step 0 / BASELINE.md 4.1: what the box offers for building the reference's own CPU path, and which GPU it has.
for t in gfortran mpirun mpicxx cmake nvcc ncu; do printf "%s: " $t; command -v $t || echo absent; done
printf "Epetra_CrsMatrix.h: "; find /usr /opt -name Epetra_CrsMatrix.h 2>/dev/null | head -1 | grep . || echo absent
printf "hdf5.h: "; find /usr /opt -name hdf5.h 2>/dev/null | head -1 | grep . || echo absent
nproc; free -g | sed -n 2p
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv,noheader 2>/dev/null || echo "no GPU visible"
