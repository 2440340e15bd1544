for g in 90x80x32 360x160x32; do
GRID=$g python tools/jac_variants.py 2>&1 | tail -1
for f in iemic_b200/lib/variants/*.so; do GRID=$g IEMIC_B200_LIB=$PWD/$f python tools/jac_variants.py 2>&1 | tail -1; done
done
