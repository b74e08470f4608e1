timeout 600 python -m pytest tests/test_tile_segments_gpu.py -x -q 2>&1 | tail -30
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for s in 1 0; do for w in 1 8; do MGS_TILE_SEGMENTS=$s python scripts/map_rate.py $w | tail -1; done; done
