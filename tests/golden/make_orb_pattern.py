import numpy as np, cv2
orb = cv2.ORB_create()
S=101; c=50
R=19
def desc_bits(img, angle=0.0):
    kp=[cv2.KeyPoint(x=float(c),y=float(c),size=20,angle=angle)]
    k,d = orb.compute(img,kp)
    assert len(k)==1
    return np.unpackbits(d[0],bitorder='little')
R1=np.zeros((256,2*R+1,2*R+1),bool); R0=np.zeros_like(R1)
for dy in range(-R,R+1):
    for dx in range(-R,R+1):
        im=np.zeros((S,S),np.uint8); im[c+dy,c+dx]=255
        R1[:,dy+R,dx+R]=desc_bits(im)
        im=np.full((S,S),255,np.uint8); im[c+dy,c+dx]=0
        R0[:,dy+R,dx+R]=desc_bits(im)
pat=np.zeros((512,2),np.int32)
bad=0
for i in range(256):
    U=R1[i]|R0[i]
    ys,xs=np.nonzero(U)
    x0,x1,y0,y1=xs.min(),xs.max(),ys.min(),ys.max()
    # squares 7x7 at corners of bbox
    if U[y0,x0] and U[y1,x1]: corners=[(x0,y0),(x1-6,y1-6)]
    else: corners=[(x1-6,y0),(x0,y1-6)]
    cand=[(ax+3-R, ay+3-R) for ax,ay in corners]
    # p1 = centre that is in R1 with bit=1 at q=p1 ; p0 in R0
    p1=[p for p in set(cand) if R1[i,p[1]+R,p[0]+R]]
    p0=[p for p in set(cand) if R0[i,p[1]+R,p[0]+R]]
    if len(p1)!=1 or len(p0)!=1: bad+=1; print('ambig',i,cand,p1,p0); continue
    pat[2*i]=p0[0]; pat[2*i+1]=p1[0]
print('bad',bad)
print(pat[:8].ravel())
print(pat.min(),pat.max())
np.save('/tmp/n3/pattern.npy',pat)
