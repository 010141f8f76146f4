for v in "" sc fo; do
  if [ -z "$v" ]; then export GNNCCA_LIB=; unset GNNCCA_LIB; else export GNNCCA_LIB=gnn-cca_b200/variants/libgnncca_$v.so; fi
  echo "=== variant ${v:-default}"
  python tools/build_stages.py 2>&1 | grep "wall: \|build_graphs device"
  python tools/build_stages.py --cams 8 --dets 64 --frames 64 --reps 30 2>&1 | grep "wall: \|build_graphs device"
done
