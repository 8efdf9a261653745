python scripts/rank_time.py --world 8 --batch-rays 0 8388608 16777216 33554432 67108864
B200RT_STREAMS=1 python scripts/rank_time.py --world 8 --batch-rays 0 16777216 67108864
python scripts/rank_time.py --world 4 --batch-rays 0 16777216 33554432
python scripts/rank_time.py --world 1 --batch-rays 0 33554432 50331648 100663296
