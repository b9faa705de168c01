#!/bin/bash
# usage: tools/sweep_sssp2.sh "EXTRAFLAGS|ENV ..." SLOTS   (each config: nvcc -D flags, then env assignments, separated by |)
cd "$(dirname "$0")/.."
python tools/prof_sssp.py 60000 4 1 > /dev/null 2>&1
IFS=';' read -ra CFGS <<< "$1"
for cfg in "${CFGS[@]}"; do
  flags="${cfg%%|*}"; envs="${cfg#*|}"
  rm -f vqvae_b200/csrc/build/sssp.o
  make -C vqvae_b200/csrc EXTRA="$flags" > /dev/null 2>&1
  env $envs PROF_TAG="[$flags|$envs]" timeout 150 python tools/prof_sssp.py 60000 $2 4 2>&1 | grep -E "slots=|FAILED"
done
rm -f vqvae_b200/csrc/build/sssp.o; make -C vqvae_b200/csrc > /dev/null 2>&1
