#!/bin/bash
# scratch: combine-phase restructure on LJ, persisting-L2 window on the TW-shaped graph (one GPU)
set -x
python tools/pr_tiers_sweep.py "12,6,1.25" "12,6,1.25"
export PR_N=41600000 PR_M=1470000000
python tools/pr_tiers_sweep.py "12,6,1.25"
GMC_PR_L2_WINDOW_VERBOSE=1 GMC_PR_L2_WINDOW_MB=24 python tools/pr_tiers_sweep.py "12,6,1.25"
GMC_PR_L2_WINDOW_MB=40 python tools/pr_tiers_sweep.py "12,6,1.25"
GMC_PR_L2_WINDOW_MB=64 GMC_PR_L2_WINDOW_HIT=0.6 python tools/pr_tiers_sweep.py "12,6,1.25"
