for g in 90x80x32 90x160x32 180x160x32; do
GRID=$g python tools/jac_variants.py 2>&1 | tail -1 | cut -c1-40
GRID=$g IEMIC_SPMV_TILE=0 python tools/jac_variants.py 2>&1 | tail -1 | cut -c1-40
done
