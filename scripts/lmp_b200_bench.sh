#!/bin/bash
# lmp_b200 -sf b200 on bench/in.lj and bench/in.eam scaled to 2 048 000 atoms (x=y=z=4 of the
# 32000-atom case), loop time as printed by the reference program's own Finish.
set -u
L=lammps-kokkos-port_b200/lammps_styles/_build/lmp_b200
mkdir -p gpurun_out
cat > /tmp/in.lj <<'E2'
units lj
atom_style atomic
lattice fcc 0.8442
region box block 0 80 0 80 0 80
create_box 1 box
create_atoms 1 box
mass 1 1.0
velocity all create 1.44 87287 loop geom
pair_style lj/cut 2.5
pair_coeff 1 1 1.0 1.0 2.5
neighbor 0.3 bin
neigh_modify delay 0 every 20 check no
fix 1 all nve
run 100
run 1000
E2
cat > /tmp/in.eam <<E2
units metal
atom_style atomic
lattice fcc 3.615
region box block 0 80 0 80 0 80
create_box 1 box
create_atoms 1 box
pair_style eam
pair_coeff 1 1 $PWD/oracle/_ref/potentials/Cu_u3.eam
velocity all create 1600.0 376847 loop geom
neighbor 1.0 bin
neigh_modify every 1 delay 5 check yes
fix 1 all nve
timestep 0.005
thermo 50
run 100
run 500
E2
$L -sf b200 -in /tmp/in.lj -log none > gpurun_out/lmp_b200_lj_2M.log 2>&1; echo "lj rc=$?"
grep -E "Loop time|Performance|^ +[0-9]+ +[0-9.-]+ " gpurun_out/lmp_b200_lj_2M.log
$L -sf b200 -in /tmp/in.eam -log none > gpurun_out/lmp_b200_eam_2M.log 2>&1; echo "eam rc=$?"
grep -E "Loop time|Performance|^ +[0-9]+ +[0-9.-]+ " gpurun_out/lmp_b200_eam_2M.log
