for b in 128 256 8192; do timeout 100 python scripts/time_tmem.py $b 2>&1 | head -1; done
