python scripts/profile_step.py 1000000 3 0 720 720 15 1 | sed 's/.*band.: \([0-9.]*\).*/headline band(15) order  \1/'
FOREMAST_GPU_NO_ORDER=1 python scripts/profile_step.py 1000000 3 0 720 720 15 1 | sed 's/.*band.: \([0-9.]*\).*/headline band(15) plain  \1/'
python scripts/profile_step.py 1000000 3 0 720 720 13 0 | sed 's/.*band.: \([0-9.]*\).*/band(13) order  \1/'
python scripts/profile_step.py 1000000 3 0 720 720 7 0 | sed 's/.*band.: \([0-9.]*\).*/band(7) order  \1/'
python scripts/profile_step.py 1000000 3 0 720 720 1 0 | sed 's/.*band.: \([0-9.]*\).*/band(1) order  \1/'
FOREMAST_GPU_NO_ORDER=1 python scripts/profile_step.py 1000000 3 0 720 720 1 0 | sed 's/.*band.: \([0-9.]*\).*/band(1) plain  \1/'
python scripts/profile_step.py 80000 6 1380 0 60 12 0 | sed 's/.*band.: \([0-9.]*\).*/cfg2-like order  \1/'
FOREMAST_GPU_NO_ORDER=1 python scripts/profile_step.py 80000 6 1380 0 60 12 0 | sed 's/.*band.: \([0-9.]*\).*/cfg2-like plain  \1/'
python scripts/ragged_time.py 200000
