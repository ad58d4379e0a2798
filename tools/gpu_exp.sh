P=reinforcement-learning_b200
python tools/phase_profile.py 160 40 2>&1 | head -1
for v in m2080 m2088 m2220 m4094; do
echo "=== $v"
GGL_LIB_PATH=$P/libgigaenv_var_$v.so timeout 300 python tools/phase_profile.py 160 40 2>&1 | head -1
done
