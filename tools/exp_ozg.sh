for kc in 32 64 128; do echo "== KC=$kc"; GPK_OZG_KC=$kc python tools/fit_digits.py 8192,16384 1e-2 0 0,6 | grep "digits=6"; done
echo "== KC=64 noise 1e-4"; GPK_OZG_KC=64 python tools/fit_digits.py 8192 1e-4 0 0,6 | grep "digits=6"
