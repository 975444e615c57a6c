"""CPU model of one pipeline stage of the cube sweep (R = 2): TMA box copies with SWIZZLE_64B into
the StageLayout of march_dev.cuh, then every lane's shared-memory quad loads (LaneK) -- checks that
each lane reads the elements it should and counts LDS.128 bank conflicts per 8-lane phase.
Mirrors the C++ formulas by hand; used to check layout experiments (-DMARCH_JL_SWIZZLE) without a GPU.
usage: python tools/emulate_stage_layout.py [P] [R]   (R = 2: 125-point cube, R = 4: 25-point star)
       python tools/emulate_stage_layout.py edge      (the -DMARCH_EDGE_HI lane numbering)"""
import sys
import numpy as np

R, HW = 2, 2  # set by main()


def layout(P, jlx):
    L = {}
    L["C_PL"], L["I_PL"], L["J_PL"], L["X_PL"] = 512, 64 * HW, 64 * R, 8 * R * HW
    L["JL_ROWS"] = 4 if jlx else R
    L["JL_PL"] = 64 * L["JL_ROWS"]
    L["C0"], L["C1"] = 0, P * 512
    L["IL"] = 2 * P * 512
    L["IR"] = L["IL"] + P * L["I_PL"]
    L["JL0"] = L["IR"] + P * L["I_PL"]
    L["JL1"] = L["JL0"] + P * L["JL_PL"]
    L["JH0"] = L["JL1"] + P * L["JL_PL"]
    L["JH1"] = L["JH0"] + P * L["J_PL"]
    xs = max(P * L["X_PL"], 128)
    L["XLL"] = L["JH1"] + P * L["J_PL"]
    L["XRL"], L["XLH"], L["XRH"] = L["XLL"] + xs, L["XLL"] + 2 * xs, L["XLL"] + 3 * xs
    L["BYTES"] = (L["XRH"] + xs + 511) // 512 * 512
    return L


def tma(smem, dst, src, box, coord, swizzle):
    """src[b][k][j][i] (ids), box (bi, bj, bk, bb), coord (ci, cj, ck, cb) -> dense box at dst"""
    bi, bj, bk, bb = box
    ci, cj, ck, cb = coord
    for b in range(bb):
        for k in range(bk):
            for j in range(bj):
                for i in range(bi):
                    a = dst + (((b * bk + k) * bj + j) * bi + i) * 8
                    if swizzle:
                        a ^= ((a >> 7) & 3) << 4
                    smem[a // 8] = src[cb + b, ck + k, cj + j, ci + i]


EDGE_HI = False  # -DMARCH_EDGE_HI: edge lanes 16 + bb * 8 + j, inner lanes (j >> 2) * 8 + (j & 3) * 2 + bb


def lane_fields(lane):
    if EDGE_HI:
        if lane >> 4:
            bb = (lane >> 3) & 1
            return bb, lane & 7, bb
        bb = lane & 1
        return bb, ((lane >> 3) & 1) * 4 + ((lane >> 1) & 3), 1 - bb
    return lane >> 4, (lane >> 1) & 7, lane & 1


def lane_of(bb, j, ih):
    if EDGE_HI:
        return 16 + bb * 8 + j if ih == bb else (j >> 2) * 8 + (j & 3) * 2 + bb
    return bb * 16 + j * 2 + ih


def shuffle_sources(lane):
    """(srcL, srcR) as LaneK::init computes them"""
    bb, j, ih = lane_fields(lane)
    if EDGE_HI:
        src_l = lane_of(bb, j, 0) if ih else (lane_of(0, j, 1) if bb else lane)
        src_r = (lane if bb else lane_of(1, j, 0)) if ih else lane_of(bb, j, 1)
    else:
        src_l = lane - 1 if ih else (lane - 15 if bb else lane)
        src_r = (lane if bb else lane + 15) if ih else lane + 1
    return src_l, src_r


def check_shuffles():
    """the left / right source lane owns the quad next to this lane's in the 16-wide row, or is the
    lane itself on the footprint edge (where the i-face value is used instead)"""
    bad = 0
    for lane in range(32):
        bb, j, ih = lane_fields(lane)
        pos = bb * 2 + ih  # quad index 0..3 along i
        for src, want in zip(shuffle_sources(lane), (pos - 1, pos + 1)):
            sb, sj, si = lane_fields(src)
            if 0 <= want <= 3:
                bad += int((sj, sb * 2 + si) != (j, want))
            else:
                bad += int(src != lane)
    edge = [lane for lane in range(32) if lane_fields(lane)[2] == lane_fields(lane)[0]]
    return bad, edge


def lanek(L, lane, jlx):
    bb, j, ih = lane_fields(lane)
    row0, row1, pitch, jx = {}, {}, {}, {}
    for d in range(-R, R + 1):
        jj, a = j + d, d + R
        jx[a] = 0
        if jlx and jj < 0:
            rr = jj + L["JL_ROWS"]
            base, x0 = (L["JL1"] if bb else L["JL0"]) + rr * 64, (rr >> 1) & 3
            row0[a], row1[a] = base + (((2 * ih) ^ x0) << 4), base + (((2 * ih + 1) ^ x0) << 4)
            pitch[a], jx[a] = L["JL_PL"], 0x20
        elif jlx and R == 4 and jj > 7:  # JHX: the +j face of the 25-point star, swizzled as well
            rr = jj - 8
            base, x0 = (L["JH1"] if bb else L["JH0"]) + rr * 64, (rr >> 1) & 3
            row0[a], row1[a] = base + (((2 * ih) ^ x0) << 4), base + (((2 * ih + 1) ^ x0) << 4)
            pitch[a], jx[a] = L["J_PL"], 0x20
        elif 0 <= jj < 8:
            base, sw = (L["C1"] if bb else L["C0"]) + jj * 64, (jj >> 1) & 3
            row0[a], row1[a] = base + (((2 * ih) ^ sw) << 4), base + (((2 * ih + 1) ^ sw) << 4)
            pitch[a] = 512
        else:
            low = jj < 0
            rr = jj + R if low else jj - 8
            base = ((L["JL1"] if bb else L["JL0"]) if low else (L["JH1"] if bb else L["JH0"])) + rr * 64
            row0[a], row1[a], pitch[a] = base + ih * 32, base + ih * 32 + 16, L["J_PL"]
    return bb, j, ih, row0, row1, pitch, jx


def run(P, jlx):
    L = layout(P, jlx)
    # bricks: 0,1 = the pair; 2,3 = their -j neighbours; 4,5 = their +j neighbours
    src = np.arange(6 * 8 * 8 * 8, dtype=np.int64).reshape(6, 8, 8, 8)
    smem = np.full(L["BYTES"] // 8, -1, dtype=np.int64)
    ck = 0
    tma(smem, L["C0"], src, (8, 8, P, 2), (0, 0, ck, 0), True)
    if jlx:
        tma(smem, L["JL0"], src, (8, 4, P, 2), (0, 4, ck, 2), True)
    else:
        tma(smem, L["JL0"], src, (8, R, P, 2), (0, 8 - R, ck, 2), False)
    tma(smem, L["JH0"], src, (8, R, P, 2), (0, 0, ck, 4), jlx and R == 4)
    wrong = conflicts = loads = 0
    for z in range(P):
        for a in range(2 * R + 1):
            for q in range(2):
                addrs = []
                for lane in range(32):
                    bb, j, ih, row0, row1, pitch, jx = lanek(L, lane, jlx)
                    off = (row1 if q else row0)[a] + z * pitch[a]
                    if jlx and (a < R or R == 4) and (z & 1):
                        off ^= jx[a]
                    addrs.append(off)
                    jj = j + a - R
                    brick, row = (bb, jj) if 0 <= jj < 8 else ((2 + bb, jj + 8) if jj < 0 else (4 + bb, jj - 8))
                    exp = src[brick, ck + z, row, 4 * ih + 2 * q: 4 * ih + 2 * q + 2]
                    got = smem[off // 8: off // 8 + 2]
                    wrong += int(not np.array_equal(exp, got))
                for ph in range(4):
                    g = [(x >> 4) & 7 for x in addrs[8 * ph: 8 * ph + 8]]
                    conflicts += max(g.count(v) for v in set(g)) - 1  # extra wavefronts of this phase
                loads += 1
    return wrong, conflicts, loads, L["BYTES"]


def halo_wavefronts(P=4):
    """wavefronts of the five i-face LDS.128 of one cube plane: per 8-lane phase, the largest number of
    DISTINCT 16-byte addresses that fall on one granule; a phase without an active lane costs nothing.
    Product numbering: every lane loads (pairs broadcast); EDGE_HI: only the edge lanes do."""
    L = layout(P, False)
    total = 0
    for d in range(-R, R + 1):
        addrs = []
        for lane in range(32):
            bb, j, ih = lane_fields(lane)
            if EDGE_HI and ih != bb:
                addrs.append(None)
                continue
            jj = j + d
            if 0 <= jj < 8:
                a = (L["IR"] if bb else L["IL"]) + jj * HW * 8
            elif jj < 0:
                a = (L["XRL"] if bb else L["XLL"]) + (jj + R) * HW * 8
            else:
                a = (L["XRH"] if bb else L["XLH"]) + (jj - 8) * HW * 8
            addrs.append(a)
        for ph in range(4):
            act = set(x >> 4 for x in addrs[8 * ph: 8 * ph + 8] if x is not None)
            if act:
                g = [x & 7 for x in act]
                total += max(g.count(v) for v in set(g))
    return total


def main_edge():
    global EDGE_HI
    for flag in (False, True):
        EDGE_HI = flag
        bad, edge = check_shuffles()
        quarters = sorted(set(x >> 3 for x in edge))
        print("   i-face loads: %d wavefronts per plane" % halo_wavefronts())
        w, c, n, _ = run(4, True)
        print("   with the swizzled -j face: wrong quads %d, %d extra wavefronts over %d LDS.128" % (w, c, n))
        w, c, n, _ = run(4, False)
        print("edge lanes high %-5s: wrong shuffle sources %d, wrong quads %d, %d extra wavefronts over %d LDS.128, "
              "edge lanes sit in quarter-warps %s" % (flag, bad, w, c, n, quarters))
    EDGE_HI = False


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "edge":
        main_edge()
        sys.exit(0)
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    R = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    HW = 4 if R > 2 else 2
    for jlx in (False, True):
        w, c, n, b = run(P, jlx)
        print("P=%d swizzled -j face %-5s: wrong quads %d, %d extra wavefronts over %d LDS.128 (%d wavefronts ideal), "
              "stage %d bytes" % (P, jlx, w, c, n, 4 * n, b))
