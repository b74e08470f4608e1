timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in 1 8; do python scripts/map_rate.py $w | tail -1; done
python scripts/prof_map.py 2 events 2>/dev/null | grep -i "blend\|sum per\|ok"
