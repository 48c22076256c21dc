# Holt-Winters kernels across periods: tensor-memory kernel (where fm_hw_tm_lanes allows it) against the shared-memory one,
# ~2 GB of samples per run.  usage: bash scripts/hw_periods.sh [periods...]
for P in ${@:-1440 1024 1000 168 288 1536}; do
  H=$((P*7)); U=$((2000000000/(H*4)))
  echo "period $P hist $H units $U"
  FM_PERIOD=$P python scripts/profile_step.py $U 3 $H 0 60 16 0 | sed 's/.*band.: \([0-9.]*\).*/  default \1 ms/'
  FM_PERIOD=$P FOREMAST_GPU_HW_SMEM=1 python scripts/profile_step.py $U 3 $H 0 60 16 0 | sed 's/.*band.: \([0-9.]*\).*/  k_hw    \1 ms/'
done
