#!/bin/bash
# Counts of the SASS mnemonics that show what the kernels of libsart_b200.so are made of, per kernel family
# (cuobjdump runs here, without a GPU): UBLKCP = bulk-async (TMA) copy, UTMALDG = TMA tensor copy (cp.async.bulk.tensor), SYNCS = mbarrier, STAS = st.async into a peer
# CTA's shared memory, LDG.128 = 128-bit global loads (LDG.E[...].128), LDS.128 = 128-bit shared loads, STL/LDL = local-memory spills.
SO=${1:-mpi-cuda-sartsolver_b200/libsart_b200.so}
cuobjdump -sass "$SO" | awk '
/Function :/ { name=$3; sub(/^_ZN4sart[0-9]*/, "", name); fam=name; sub(/I[Lj].*$/, "", fam); sub(/E[vP].*$/, "", fam) }
/EF_CUDA_SM/ { arch=$0 }
{ for (i = 1; i <= n; i++) if ($0 ~ pat[i]) count[fam, key[i]]++ ; if (fam != "") seen[fam] = 1 }
BEGIN { n = split("UBLKCP UTMALDG SYNCS STAS LDG.128 LDS.128 FFMA DADD STL LDL ATOMG MEMBAR", key, " ");
        for (i = 1; i <= n; i++) pat[i] = "[ \t]" key[i]; pat[5] = "LDG\\.E[.A-Z]*\\.128"; pat[6] = "LDS\\.128" }
END {
  printf "%-34s", "kernel family (all instances)"; for (i = 1; i <= n; i++) printf "%10s", key[i]; printf "\n";
  for (f in seen) { printf "%-34s", f; for (i = 1; i <= n; i++) printf "%10d", count[f, key[i]]; printf "\n" }
}' | sort
cuobjdump -lelf "$SO" | head -3
