for v in "" _nosqrt _noex2 _nomma _rsq _nomufu; do
  echo "variant $v: $(OB200_LIB=$PWD/metaopt_b200/libob200$v.so python tools/bound_sweep.py 2>&1 | tail -1)"
done
