"""Command-line / config-file options of the reference (nerf_shared/config_parser.py:2-114).

Same flag names, types and defaults, declared from a table; `key = value` config files are read
with `--config` as in the reference.  Uses configargparse when it is installed and a small
argparse front-end with the same file syntax otherwise (the package is not always available)."""
import argparse

try:
    import configargparse as _cap
except Exception:  # pragma: no cover
    _cap = None



def _str2bool(v):
    """`tensorboard = False` in a config file must mean False (argparse's type=bool maps every non-empty string,
    including "False", to True: config_parser.py:113 of the reference has that trap)."""
    if isinstance(v, bool):
        return v
    if str(v).strip().lower() in ("1", "true", "yes", "y", "on"):
        return True
    if str(v).strip().lower() in ("0", "false", "no", "n", "off", ""):
        return False
    raise argparse.ArgumentTypeError("expected a boolean, got %r" % (v,))


# (flag, kwargs)
_OPTIONS = [
    # store_true as in the reference (config_parser.py:14): configs without `training = True` (e.g. fern.txt) do not train
    ("--training", dict(action="store_true", help="run the training loop of main.py")),
    ("--expname", dict(type=str, help="experiment name")),
    ("--basedir", dict(type=str, default="./logs/", help="where to store ckpts and logs")),
    ("--datadir", dict(type=str, default="./data/llff/fern", help="input data directory")),
    ("--netdepth", dict(type=int, default=8)), ("--netwidth", dict(type=int, default=256)),
    ("--netdepth_fine", dict(type=int, default=8)), ("--netwidth_fine", dict(type=int, default=256)),
    ("--N_rand", dict(type=int, default=32 * 32 * 4, help="rays per gradient step")),
    ("--lrate", dict(type=float, default=5e-4)), ("--lrate_decay", dict(type=int, default=250)),
    ("--chunk", dict(type=int, default=1024 * 32)), ("--netchunk", dict(type=int, default=1024 * 64)),
    ("--no_batching", dict(action="store_true")), ("--no_reload", dict(action="store_true")),
    ("--ft_path", dict(type=str, default=None)),
    ("--N_samples", dict(type=int, default=64)), ("--N_importance", dict(type=int, default=0)),
    ("--perturb", dict(type=float, default=1.0)), ("--use_viewdirs", dict(action="store_true")),
    ("--i_embed", dict(type=int, default=0)), ("--multires", dict(type=int, default=10)),
    ("--multires_views", dict(type=int, default=4)), ("--raw_noise_std", dict(type=float, default=0.0)),
    ("--render_only", dict(action="store_true")), ("--render_test", dict(action="store_true")),
    ("--render_factor", dict(type=int, default=0)),
    ("--precrop_iters", dict(type=int, default=0)), ("--precrop_frac", dict(type=float, default=0.5)),
    ("--dataset_type", dict(type=str, default="llff")), ("--testskip", dict(type=int, default=8)),
    ("--shape", dict(type=str, default="greek")), ("--white_bkgd", dict(action="store_true")),
    ("--half_res", dict(action="store_true")),
    ("--factor", dict(type=int, default=8)), ("--no_ndc", dict(action="store_true")),
    ("--lindisp", dict(action="store_true")), ("--spherify", dict(action="store_true")),
    ("--llffhold", dict(type=int, default=8)),
    ("--i_print", dict(type=int, default=100)), ("--i_img", dict(type=int, default=500)),
    ("--i_weights", dict(type=int, default=10000)), ("--i_testset", dict(type=int, default=50000)),
    ("--i_video", dict(type=int, default=50000)),
    ("--tensorboard", dict(type=_str2bool, default=False)),
]


class _FileArgParser(argparse.ArgumentParser):
    """argparse + `--config file` with `key = value` lines (command line wins over the file)."""

    def parse_known_args(self, args=None, namespace=None):
        import sys
        argv = list(sys.argv[1:] if args is None else args)
        pre = argparse.ArgumentParser(add_help=False)
        pre.add_argument("--config")
        known, _ = pre.parse_known_args(argv)
        file_args = []
        if known.config:
            flags = {a.option_strings[0]: a for a in self._actions if a.option_strings}
            with open(known.config) as f:
                for line in f:
                    line = line.split("#")[0].strip()
                    if not line or "=" not in line:
                        continue
                    k, v = (t.strip() for t in line.split("=", 1))
                    act = flags.get("--" + k)
                    if act is None:
                        continue          # undeclared keys are ignored (e.g. i_test in tb_lego.txt)
                    if isinstance(act, argparse._StoreTrueAction):
                        if v.lower() in ("true", "1", "yes"):
                            file_args.append("--" + k)
                    else:
                        file_args += ["--" + k, v]
        return super().parse_known_args(file_args + argv, namespace)


def config_parser():
    if _cap is not None:
        parser = _cap.ArgumentParser()
        parser.add_argument("--config", is_config_file=True, help="config file path")
    else:
        parser = _FileArgParser()
        parser.add_argument("--config", type=str, default=None, help="config file path")
    for flag, kw in _OPTIONS:
        parser.add_argument(flag, **kw)
    return parser
