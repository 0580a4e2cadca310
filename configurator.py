"""Poor man's configurator, nanoGPT-compatible (SURVEY.md 8b "CLI").

Usage, exactly as the reference's notebook / k8s jobs call it
(/root/reference/notebooks/colab_nanoGPT_companion.ipynb:70-79):

    python train.py config/train_shakespeare_char.py --batch_size=32 --compile=False

Meant to be run with ``exec(open('configurator.py').read())`` from a script whose settings are
module globals: a bare argument is a config file that is exec'd over those globals, a
``--key=value`` argument overrides one global, must name an existing key and must keep its type.
"""
import sys
from ast import literal_eval

for _arg in sys.argv[1:]:
    if "=" not in _arg:
        if _arg.startswith("--"):
            raise ValueError(f"flag {_arg!r} needs a value: use --key=value")
        print(f"Overriding config with {_arg}:")
        with open(_arg) as _f:
            _src = _f.read()
        print(_src)
        exec(_src)
    else:
        if not _arg.startswith("--"):
            raise ValueError(f"override {_arg!r} must look like --key=value")
        _key, _val = _arg[2:].split("=", 1)
        if _key not in globals():
            raise ValueError(f"Unknown config key: {_key}")
        try:
            _new = literal_eval(_val)
        except (SyntaxError, ValueError):
            _new = _val  # plain string
        if type(_new) != type(globals()[_key]):  # noqa: E721  (bool must not pass for int)
            raise ValueError(f"--{_key}: expected {type(globals()[_key]).__name__}, got {type(_new).__name__}")
        print(f"Overriding: {_key} = {_new}")
        globals()[_key] = _new
