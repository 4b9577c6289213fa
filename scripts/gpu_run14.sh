timeout 120 python scripts/debug_times.py
