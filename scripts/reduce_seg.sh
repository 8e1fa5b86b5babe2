for seg in 256 512 1024; do echo "seg=$seg"; H2B_MSM_REDUCE_SEG=$seg python scripts/knob_ab.py msm_canon_once 1 1 20,21,24 10 2>&1 | grep -E "2\^" | awk 'NR%6==1' | cut -c1-230; done
