set -e
cd 3dgs_b200/csrc
for mb in 3 4 5 6; do
  rm -f build/batch.o; make -s EXTRA="-DGSX_BATCH_MINB=$mb" >/dev/null 2>&1
  cd ../..; echo "MINB=$mb"; python profiles/stage_times.py 1 3 6 | cut -c1-400; python profiles/stage_times.py 8 2 2 | cut -c1-400; cd 3dgs_b200/csrc
done
rm -f build/batch.o; make -s >/dev/null 2>&1
