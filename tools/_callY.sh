export PROF_ONLY_WW=1
echo "== tree"; python tools/prof_train.py 20 100 5
for v in A B C; do echo "== $v"; SRNN_LIB=$PWD/build/variants/train_$v.so python tools/prof_train.py 20 100 5; done
