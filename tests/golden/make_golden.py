"""Generates tests/golden/*.npz from the reference's own fixtures and code (run in the build
container, where /root/reference is mounted; the GPU box only reads the .npz files).

  fixture_siso.npz  tools/sig80211GenMultipleSiso_1x1_0.bin  -> reference signal -> demod -> decode
  fixture_mimo.npz  tools/sig80211GenMultipleMimo_2x2_{0,1}.bin -> signal2 -> demod2 -> decode
  frame94.npy       the 94-byte MPDU of tools/debug.txt:5 both fixtures carry

Every array is an OUTPUT of the unmodified reference blocks (oracle/_ref) or a byte-for-byte
view of the reference's data files; nothing here is reference source.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import refgr  # noqa: E402

REF = "/root/reference/tools"


def run(files, sync_idx, name):
    iq = [np.fromfile(os.path.join(REF, f), np.complex64) for f in files]
    sig = refgr.run_signal(iq, sync_idx)
    tags = sig[0].tags()
    assert len(tags) == 1
    off, t = tags[0]
    nsamp = t["nsamp"]
    post = [s.data()[off:off + nsamp] for s in sig]
    llr_stream = refgr.run_demod(sig)
    pk = refgr.split_llr_stream(llr_stream)
    assert len(pk) == 1
    msgs = refgr.run_decode(llr_stream)
    assert len(msgs) == 1
    d = pk[0]
    np.savez_compressed(os.path.join(HERE, name), post=np.stack(post), l_mcs=t["mcs"], l_len=t["len"], nsamp=nsamp,
                        chan=t["chan"], cfo=t["cfo"], llr=d["llr"],
                        tags=np.array([d[k] for k in ("format", "mcs", "len", "cr", "ampdu", "trellis", "total")], np.int64),
                        msg=np.frombuffer(msgs[0], np.uint8))
    return msgs[0]


def main():
    txt = [ln for ln in open(os.path.join(REF, "debug.txt")).read().splitlines() if ln.startswith("mac packet 94 ")][0]
    frame = bytes.fromhex(txt.split()[3])
    assert len(frame) == 94, len(frame)
    np.save(os.path.join(HERE, "frame94.npy"), np.frombuffer(frame, np.uint8))
    m1 = run(["sig80211GenMultipleSiso_1x1_0.bin"], 1200 + 160 + 16, "fixture_siso.npz")
    m2 = run(["sig80211GenMultipleMimo_2x2_0.bin", "sig80211GenMultipleMimo_2x2_1.bin"], 1200 + 160 + 16, "fixture_mimo.npz")
    assert m1[3:-1] == frame and m2[3:-1] == frame
    print("golden written:", m1[:3].hex(), m2[:3].hex(), m1[-1], m2[-1])


if __name__ == "__main__":
    main()
