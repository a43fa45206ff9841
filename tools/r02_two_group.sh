python tools/two_group.py 1024 32 400 1
GPB_CTAS_PER_SM_RT=2 GPB_GRID_CAP=1 python tools/two_group.py 1024 32 400 1
for d in 0 100 200 300; do GPB_CTAS_PER_SM_RT=2 GPB_GRID_CAP=1 python tools/two_group.py 1024 32 400 2 $d; done
for d in 0 150; do python tools/two_group.py 1024 32 400 2 $d; done
for d in 0 100; do GPB_CTAS_PER_SM_RT=2 GPB_GRID_CAP=1 python tools/two_group.py 1024 32 400 4 $d; done
GPB_CTAS_PER_SM_RT=2 python tools/two_group.py 1024 32 400 2 150
