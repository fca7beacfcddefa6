"""Generates the static schedule tables of csrc/spconv_octet.cu (printed as C initialisers).

A k3 convolution on COMPLETE octets (children 8p+o of parent p, o = ox + 2 oy + 4 oz) equals a k3 convolution on the
parents with block-sparse weights: child o' of the neighbour parent p + D contributes to child o of p through
W[d], d = 2 D + o' - o, whenever d is in {-1,0,1}^3.  A tile (128 parents) is processed in 16 stages of 4 units (+ 4 half-filled stages for a residual input);
a unit = (parent offset D, source child o') = one gathered [128 x Cin] A operand.
  c_unit[S][u]       = jD | o' << 8                       (jD = (Dx+1) + 3 (Dy+1) + 9 (Dz+1))
  c_mma[S][u][mw]    = valid | n2 << 1 | q << 2 | col << 8   issuer mw owns the output children (ox, oy, oz) with
                       oy + 2 oz = mw: col = first output child, n2 = both ox (N = 2 cout), q = 26 - jd of the
                       first panel (panels are stored in reversed offset order so that ox -> ox + 1 is q -> q + 1)
"""
import itertools


def jd(d):
    return (d[0] + 1) + 3 * (d[1] + 1) + 9 * (d[2] + 1)


def children(D):
    """source children o' of neighbour parent p + D that touch parent p"""
    opts = [((0, 1) if D[a] == 0 else ((0,) if D[a] == 1 else (1,))) for a in range(3)]
    return [(x, y, z) for z in opts[2] for y in opts[1] for x in opts[0]]


def oidx(o):
    return o[0] + 2 * o[1] + 4 * o[2]


def stages():
    S = []
    Ds = list(itertools.product((-1, 0, 1), repeat=3))
    by = lambda pred: [D for D in Ds if pred(D)]
    nz = lambda D: sum(1 for v in D if v)
    centre = [(0, 0, 0)]
    yzface = by(lambda D: D[0] == 0 and nz(D) == 1)
    yzedge = by(lambda D: D[0] == 0 and nz(D) == 2)
    xface = by(lambda D: D[0] != 0 and nz(D) == 1)
    xedge = by(lambda D: D[0] != 0 and nz(D) == 2)
    corner = by(lambda D: nz(D) == 3)
    units = []
    for grp in (centre, yzface, yzedge, xface, xedge, corner):
        for D in grp:
            ch = children(D)
            units += [(D, o) for o in sorted(ch, key=oidx)]
    assert len(units) == 64
    return [units[4 * s: 4 * s + 4] for s in range(16)]


def main():
    st = stages()
    unit_tab, mma_tab = [], []
    blocks = pairs = 0
    for S, us in enumerate(st):
        if S < 8:      # paired stages: units (0,1) and (2,3) are x-siblings of one D
            for a, b in ((0, 1), (2, 3)):
                assert us[a][0] == us[b][0] and us[a][0][0] == 0 and oidx(us[b][1]) == oidx(us[a][1]) + 1
        row_u, row_m = [], []
        for (D, o) in us:
            row_u.append(jd(D) | (oidx(o) << 8))
            ms = []
            for mw in range(4):
                oy, oz = mw & 1, mw >> 1
                dy, dz = 2 * D[1] + o[1] - oy, 2 * D[2] + o[2] - oz
                if abs(dy) > 1 or abs(dz) > 1:
                    ms.append(0)
                    continue
                if D[0] == 0:
                    d0 = (o[0], dy, dz)              # output ox = 0; ox = 1 uses d0 - (1,0,0): the next panel
                    q, col, n2 = 26 - jd(d0), 2 * mw, 1
                    assert 26 - jd((o[0] - 1, dy, dz)) == q + 1
                    blocks += 2
                else:
                    ox = 1 if D[0] == 1 else 0
                    dx = 2 * D[0] + o[0] - ox
                    assert abs(dx) == 1
                    q, col, n2 = 26 - jd((dx, dy, dz)), 2 * mw + ox, 0
                    blocks += 1
                pairs += 1
                ms.append(1 | (n2 << 1) | (q << 2) | (col << 8))
            row_m.append(ms)
        unit_tab.append(row_u)
        mma_tab.append(row_m)
    assert blocks == 216 and pairs == 144, (blocks, pairs)
    # stages 16..19: the residual rows of the tile (children 2r, 2r+1 of the parent itself, read from the residual
    # tensor) times the identity panel (slot 27): only issuer r has work in stage 16 + r; units 2, 3 stay empty
    for r in range(4):
        unit_tab.append([13 | ((2 * r) << 8), 13 | ((2 * r + 1) << 8), 0xffff, 0xffff])
        row = []
        for u in range(4):
            row.append([(1 | (27 << 2) | ((2 * r + u) << 8)) if (mw == r and u < 2) else 0 for mw in range(4)])
        mma_tab.append(row)
    # first MMA of a tile for every issuer is (stage 0, unit 0): it initialises both of the issuer's column blocks
    assert all((m & 3) == 3 for m in mma_tab[0][0])
    print("__constant__ uint16_t c_unit[20][4] = {")
    for r in unit_tab:
        print("    {" + ", ".join(f"0x{v:04x}" for v in r) + "},")
    print("};")
    print("constexpr uint16_t k_mma[20][4][4] = {")
    for r in mma_tab:
        print("    {" + ", ".join("{" + ", ".join(f"0x{v:04x}" for v in ms) + "}" for ms in r) + "},")
    print("};")


if __name__ == "__main__":
    main()
