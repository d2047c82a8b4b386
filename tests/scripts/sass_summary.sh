# Per-kernel SASS evidence for the Blackwell-native claims: counts of the tensor-core / TMEM / TMA mnemonics in libckb.so
# (cuobjdump -sass; runs without a GPU).  usage: bash tests/scripts/sass_summary.sh > profiles/r2_sass_summary.txt
LIB=${1:-composable_kernel_b200/lib/libckb.so}
echo "# $LIB  ($(date -u +%F)): SASS mnemonic counts per kernel family (cuobjdump -sass, sm_100a)"
echo "# UTCHMMA/UTCIMMA = tcgen05.mma kind::f16 / kind::i8 (.2CTA = cta_group::2), LDTM/STTM = tcgen05.ld/st (TMEM), UTMALDG/UTMASTG = TMA load/store,"
echo "# UTMAPF = TMA prefetch, HMMA = mma.sync, SYNCS = mbarrier ops, UTCBAR = tcgen05.commit"
cuobjdump -sass "$LIB" 2>/dev/null | awk '
  /Function :/ { fn=$3; sub(/^_ZN[0-9]*ckb[0-9]*/,"",fn); fam=fn;
      if (fn ~ /wq_fused/) fam="wq_fused_kernel (fused prefill)"; else if (fn ~ /gemm_ss2/) fam="gemm_ss2_kernel (CTA pair)";
      else if (fn ~ /gemm_ss_/) fam="gemm_ss_kernel"; else if (fn ~ /wq_gemv2/) fam="wq_gemv2_kernel (W4 decode)";
      else if (fn ~ /wq_gemv_kernel/) fam="wq_gemv_kernel (W4 decode, single producer)"; else if (fn ~ /wq_decode/) fam="wq_decode_kernel (W8 decode, stream-K)";
      else if (fn ~ /grouped/) fam="grouped_gemm_kernel"; else if (fn ~ /wq_dequant/) fam="wq_dequant kernels (two-pass prefill, comparison only)";
      else if (fn ~ /prepack|transpose/) fam="prepack kernels"; else if (fn ~ /peer_wait/) fam="peer_wait_kernel"; else fam="other";
      nk[fam]++ }
  { for (i=1;i<=NF;i++) { op=$i;
      if (op ~ /^UTCHMMA/) { c[fam,"UTCHMMA"]++; if (op ~ /2CTA/) c[fam,"UTCHMMA.2CTA"]++ }
      else if (op ~ /^UTCIMMA/) { c[fam,"UTCIMMA"]++; if (op ~ /2CTA/) c[fam,"UTCIMMA.2CTA"]++ }
      else if (op ~ /^LDTM/) c[fam,"LDTM"]++; else if (op ~ /^STTM/) c[fam,"STTM"]++;
      else if (op ~ /^UTMALDG/) c[fam,"UTMALDG"]++; else if (op ~ /^UTMASTG/) c[fam,"UTMASTG"]++; else if (op ~ /^UTMAPF/) c[fam,"UTMAPF"]++;
      else if (op ~ /^HMMA/) c[fam,"HMMA"]++; else if (op ~ /^IMMA/) c[fam,"IMMA"]++; else if (op ~ /^SYNCS/) c[fam,"SYNCS"]++; else if (op ~ /^UTCBAR/) c[fam,"UTCBAR"]++;
      else if (op ~ /^REDG|^RED\./) c[fam,"RED"]++ } }
  END { n=split("UTCHMMA UTCHMMA.2CTA UTCIMMA UTCIMMA.2CTA UTCBAR LDTM STTM UTMALDG UTMASTG UTMAPF HMMA SYNCS RED", cols, " ");
        printf "%-58s %8s", "#kernel family", "variants"; for (i=1;i<=n;i++) printf " %12s", cols[i]; printf "\n";
        for (f in nk) { printf "%-58s %8d", f, nk[f]; for (i=1;i<=n;i++) printf " %12d", c[f,cols[i]]+0; printf "\n" } }' | sort
