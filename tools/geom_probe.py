import sys; sys.path.insert(0,'.')
import dsd2dxd_b200 as d2p
for kw in [dict(dsd_rate=4,out_rate=352800,channels=2),dict(dsd_rate=4,out_rate=176400,bits=20,channels=2),dict(dsd_rate=8,out_rate=352800,channels=6,layout="I",lsb_first=False),dict(dsd_rate=1,out_rate=352800,channels=2),dict(dsd_rate=2,out_rate=88200,filt="C",channels=2),dict(dsd_rate=4,out_rate=88200,channels=2),dict(dsd_rate=1,out_rate=88200,channels=2),dict(dsd_rate=4,out_rate=192000,bits=32,channels=2),dict(dsd_rate=8,out_rate=384000,channels=6,layout="I",lsb_first=False),dict(dsd_rate=1,out_rate=96000,channels=2)]:
    try:
        b=d2p.Batch(d2p.make_config(**kw)); b.close()
    except Exception as e: print(e)
