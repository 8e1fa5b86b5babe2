for ch in 8 4; do for ba in 0 1 2; do echo "chunks=$ch ba=$ba"; H2B_MSM_CHUNKS=$ch BA_CONFIGS=$ba:0 python scripts/ba_sweep.py 24 4 2>&1 | grep -v "^{" | cut -c1-200; done; done
