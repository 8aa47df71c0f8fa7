for vn in 5 6 7 8; do for vt in 3 5; do echo "VN=$vn VT=$vt"; PTC_VOTE_NODE=$vn PTC_VOTE_TRI=$vt timeout 300 python scripts/profile_step.py 8 fast atrium 2 | cut -c29-130; done; done
