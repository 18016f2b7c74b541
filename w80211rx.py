"""Import shim: the package directory name carries hyphens (fork-gr-ieee80211_b200), so
`import w80211rx` gives its modules under plain names: w80211rx.host, w80211rx.gen, w80211rx.gen_ppdu."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
PKG = "fork-gr-ieee80211_b200"
pkg = importlib.import_module(PKG)
gen = importlib.import_module(PKG + ".gen")
gen_ppdu = importlib.import_module(PKG + ".gen_ppdu")
host = importlib.import_module(PKG + ".host")
shard = importlib.import_module(PKG + ".shard")
