"""Golden vectors from the reference model EXACTLY AS SHIPPED (MDCT2 / IMDCT2 / IDCT_2N_native, pix2pixHD_model.py:35-40)
with --use_time_D --use_match_loss --lambda_time 10: nothing in the reference is replaced or hooked.
    python tests/golden/make_golden_model_mdct2.py -> tests/golden/model_mdct2_ref.npz
Same miniature config, batches and step loop as make_golden_model_time.py (train.py:155-184)."""
from make_golden_model_time import main

if __name__ == "__main__":
    main(shipped_mdct2=True, out_name="model_mdct2_ref.npz")
