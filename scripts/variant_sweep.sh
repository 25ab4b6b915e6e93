#!/bin/bash
# time kernel-variant builds (lammps-kokkos-port_b200/libb200md_<v>.so) on the same workloads
for v in "" _x1 _x2 _x3 _x4; do
  lib=lammps-kokkos-port_b200/libb200md$v.so
  [ -f $lib ] || continue
  for wl in lj eam; do
    B200MD_LIB=$PWD/$lib B200MD_PHASE_TIMING=1 python scripts/profile_run.py $wl 100 2>&1 | tail -1 | sed "s|^|variant '$v' |"
  done
done
lib=lammps-kokkos-port_b200/libb200md_x1.so
for lanes in 1 4; do for lk in half full; do
  B200MD_LIB=$PWD/$lib B200MD_PHASE_TIMING=1 python scripts/profile_run.py lj 100 $lanes 4 $lk 2>&1 | tail -1 | sed "s|^|x1 lanes $lanes $lk |"
done; done
B200MD_LIB=$PWD/$lib B200MD_PHASE_TIMING=1 python scripts/profile_run.py lj 100 2 4 full 2>&1 | tail -1 | sed "s|^|x1 lanes 2 full |"
B200MD_LIB=$PWD/$lib B200MD_PHASE_TIMING=1 python scripts/profile_run.py eam 100 1 2>&1 | tail -1 | sed "s|^|x1 lanes 1 half |"
B200MD_LIB=$PWD/$lib B200MD_PHASE_TIMING=1 python scripts/profile_run.py eam 100 4 2>&1 | tail -1 | sed "s|^|x1 lanes 4 half |"
