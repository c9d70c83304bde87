"""Drop-in `models` package: same module names as the reference (`from models import
model_res_sigmoid_multi as md`), classes computed by libpntf_b200.so.  Put `p-ntfields_b200/` on
sys.path in place of the reference checkout's root."""
