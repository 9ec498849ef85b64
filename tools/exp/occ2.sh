for v in c5 c4 c3 c2 c1; do echo $v; ME_B200_TRACE_LIB=tools/exp/lib_$v.so python tools/profile_step.py 6 motion 3840:2160:4:2:2 2>&1 | grep "^poc" | tail -4 | cut -c1-100; done
