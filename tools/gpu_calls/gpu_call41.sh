python tools/dbg_traj_lists.py 2>&1 | tail -14
