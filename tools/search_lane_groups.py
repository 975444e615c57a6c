"""Searches the grouping of the 32 lanes into quarter-warps for the cheapest shared-memory traffic of a cube
plane (quad loads for every row offset and both chunks, i-face loads by the edge lanes only), using the
address model of tools/emulate_stage_layout.py.  Result: DESIGN.md 4.1 (5), -DMARCH_EDGE_HI.
usage: python tools/search_lane_groups.py   (about a minute)"""
import sys, itertools
import os; sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import emulate_stage_layout as E
P=4
def addr_tables(jlx):
    L=E.layout(P,jlx)
    # per (bb,j,ih): list of quad addresses for each (d,q) at z=0 and z=1, and halo addresses per d
    tab={}
    for bb in (0,1):
        for j in range(8):
            for ih in (0,1):
                # reuse lanek with product numbering lane
                E.EDGE_HI=False
                lane=bb*16+j*2+ih
                _,_,_,row0,row1,pitch,jx=E.lanek(L,lane,jlx)
                quads=[]
                for z in (0,1):
                    for a in range(5):
                        for q in (0,1):
                            off=(row1 if q else row0)[a]+z*pitch[a]
                            if jlx and a<2 and (z&1): off^=jx[a]
                            quads.append(off)
                halo=[]
                for d in range(-2,3):
                    jj=j+d
                    if 0<=jj<8: a=(L["IR"] if bb else L["IL"])+jj*16
                    elif jj<0: a=(L["XRL"] if bb else L["XLL"])+(jj+2)*16
                    else: a=(L["XRH"] if bb else L["XLH"])+(jj-8)*16
                    halo.append(a)
                tab[(bb,j,ih)]=(quads,halo)
    return tab
def wf(addrs):
    act=set(x>>4 for x in addrs)
    g=[x&7 for x in act]
    return max(g.count(v) for v in set(g))
def cost(group,tab,edge):
    c=0
    nq=len(tab[group[0]][0])
    for i in range(nq):
        c+=wf([tab[t][0][i] for t in group])
    c=c/2.0  # two planes -> per plane
    h=0
    if edge:
        for d in range(5):
            h+=wf([tab[t][1][d] for t in group])
    return c,h
for jlx in (False,True):
    tab=addr_tables(jlx)
    edge=[(bb,j,bb) for bb in (0,1) for j in range(8)]
    non=[(bb,j,1-bb) for bb in (0,1) for j in range(8)]
    res={}
    for name,lanes,isedge in (("edge",edge,True),("inner",non,False)):
        best=None
        first=lanes[0]
        rest=lanes[1:]
        for comb in itertools.combinations(rest,7):
            g1=(first,)+comb
            g2=tuple(t for t in lanes if t not in g1)
            c1,h1=cost(g1,tab,isedge); c2,h2=cost(g2,tab,isedge)
            tot=c1+c2+h1+h2
            if best is None or tot<best[0]: best=(tot,c1+c2,h1+h2,g1,g2)
        res[name]=best
        print("jlx",jlx,name,"best total/plane",best[0],"quads",best[1],"halo",best[2]); print("   ",best[3]); print("   ",best[4])
    # product numbering for reference
    E.EDGE_HI=False
    tot=0;hal=0
    for qtr in range(4):
        g=[E.lane_fields(l) for l in range(8*qtr,8*qtr+8)]
        c,h=cost(g,tab,True); tot+=c; hal+=h
    print("jlx",jlx,"product numbering: quads",tot,"halo(all lanes load)",hal)
