"""B200-native receive hot path of gr-ieee80211 (demod / demod2 / decode).

The directory name carries hyphens, so import it with
``importlib.import_module("fork-gr-ieee80211_b200")`` or through the
``w80211rx`` shim at the repository root.  The CUDA library is loaded lazily by
``.host`` (there is no CPU fallback: a missing library raises).
"""
