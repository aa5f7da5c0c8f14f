for kb in "32,56,104,200" "20,32,56,104" "24,36,56,104" "16,24,40,104" "20,28,40,104" "24,32,48,104" "18,26,36,104"; do
  GZCUDA_PGS_CHAIN_KB=$kb python scripts/dev_pgs_knobs.py 2>&1 | tail -1 | cut -c1-60
done
for cm in 20 24; do for kb in "20,32,56,104" "20,28,40,104"; do
  GZCUDA_PGS_CHAIN_MIN=$cm GZCUDA_PGS_CHAIN_KB=$kb python scripts/dev_pgs_knobs.py 2>&1 | tail -1 | cut -c1-60
done; done
