run final PRT_POOL_RAYS=2
run final_b7 PRT_POOL_RAYS=2 PRT_LIB=$PWD/prosper_b200/libprt_v_b7.so
run final_spl8 PRT_POOL_RAYS=2 -- --samples-per-launch 8
