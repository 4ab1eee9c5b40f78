from totalsegmentator2d_b200.main import ts2d_entry_point

ts2d_entry_point()
