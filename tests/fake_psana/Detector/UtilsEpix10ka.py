"""Stub of ``Detector.UtilsEpix10ka``: the control-bit helper PsanaSource prefers when psana is importable."""
import numpy as np


class _Asic:
    def __init__(self, tr):
        self._tr = tr

    def trbit(self):
        return self._tr


class _Elem:
    """One segment's configuration: asicPixelConfigArray (rows, cols) and four ASICs with a trbit each.
    ASIC -> quadrant as psana lays them out: A0 bottom right, A1 top right, A2 top left, A3 bottom left."""
    QUADS = ((1, 1), (0, 1), (0, 0), (1, 0))

    def __init__(self, cfg_seg):
        self._cfg = cfg_seg
        rows, cols = cfg_seg.shape
        self._tr = [int(cfg_seg[r * (rows // 2), c * (cols // 2)] >> 4) & 1 for r, c in self.QUADS]

    def asicPixelConfigArray(self):
        return (self._cfg & 0x0F).astype(np.uint16)                     # bits 3,2 = gain config, bits 1,0 = mask / test

    def asics(self, i):
        return _Asic(self._tr[i])


class _Cob:
    def __init__(self, cfg):
        self._elems = [_Elem(cfg[s]) for s in range(cfg.shape[0])]

    def numberOfElements(self):
        return len(self._elems)

    def elemCfg(self, i):
        return self._elems[i]


def get_epix10ka_any_config_object(env, src):
    import psana
    det = psana.Detector(src) if isinstance(src, str) else src
    return _Cob(det._c["cfg"])


def cbits_config_epix10ka_any(cob):
    """(nseg, rows, cols): pixel gain-config bits 3,2 with the ASIC's trbit in bit 4."""
    out = []
    for i in range(cob.numberOfElements()):
        e = cob.elemCfg(i)
        cb = (np.asarray(e.asicPixelConfigArray()) & 12).astype(np.uint8)
        rows, cols = cb.shape
        for a, (r, c) in enumerate(_Elem.QUADS):
            if e.asics(a).trbit():
                cb[r * (rows // 2):(r + 1) * (rows // 2), c * (cols // 2):(c + 1) * (cols // 2)] |= 16
        out.append(cb)
    return np.stack(out)
