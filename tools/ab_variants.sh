# A/B harness for kernel variants: same box, same commands, one build of the product per variant.
#   python quest_b200/build.py --variant NAME -DQB_TILE_MAXREG=112 ...     (here, no GPU needed: lib/variants/libNAME.so)
#   bash tools/ab_variants.sh NAME1 NAME2 ...                               (on the GPU box; "" = the product itself)
# A variant suffixed with + also runs the fusion parity tests.  QUEST_B200_LIB selects the library (quest_b200/capi.py).
for spec in "" "$@"; do
  v=${spec%+}
  if [ -n "$v" ]; then export QUEST_B200_LIB=$PWD/quest_b200/lib/variants/lib$v.so; else unset QUEST_B200_LIB; fi
  if [ "$spec" != "$v" ]; then
    python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -1
  fi
  QUEST_B200_PRINT_PASS_TIMES=1 python tools/profile_tile.py --steps 1 2>&1 | tail -2 | head -1 | sed "s/^/[${v:-product}] /"
done
unset QUEST_B200_LIB
