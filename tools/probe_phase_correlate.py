"""Build-container probe: cv2.phaseCorrelate on the sigma-7 blurred demo pair against NumPy restatements of
the normalised cross-power spectrum.  The correlation surface of the blurred frames is flat over +-2 px (191 720 at the peak,
190 914 / 160 360 beside it), so float32 vs float64 spectra move the argmax by 4 px and the 5x5 centroid by more than a
pixel: a device phaseCorrelate that does not reproduce OpenCV's DFT bit for bit cannot be pinned to the reference.  The
drop-in therefore keeps the host cv2 calls for real pairs (shape_ftp._align_host)."""
import cv2, numpy as np, time
G='/root/repo/tests/golden/'
ref=cv2.imread(G+'FINAL_reference.jpg'); de=cv2.imread(G+'FINAL_E_deformed.jpg')
rg=cv2.cvtColor(ref,cv2.COLOR_BGR2GRAY).astype(np.float32); dg=cv2.cvtColor(de,cv2.COLOR_BGR2GRAY).astype(np.float32)
rb=cv2.GaussianBlur(rg,(0,0),7); db=cv2.GaussianBlur(dg,(0,0),7)
h,w=rb.shape
hann=cv2.createHanningWindow((w,h),cv2.CV_32F)
t=time.time(); shift,resp=cv2.phaseCorrelate(np.float32(rb),np.float32(db),hann); print("cv2",shift,resp,time.time()-t)
# hann formula check
wc=0.5*(1-np.cos(2*np.pi*np.arange(w)/(w-1))); wr=0.5*(1-np.cos(2*np.pi*np.arange(h)/(h-1)))
h1=np.sqrt(np.outer(wr,wc)).astype(np.float32); print("hann sqrt diff",np.abs(h1-hann).max())
h2=np.outer(wr,wc).astype(np.float32); print("hann plain diff",np.abs(h2-hann).max())
def pc(a,b,win):
    A=np.fft.fft2((a*win).astype(np.float32)); B=np.fft.fft2((b*win).astype(np.float32))
    P=A*np.conj(B); C=P/(np.abs(P)+np.finfo(np.float32).eps)
    c=np.real(np.fft.ifft2(C)).astype(np.float32)*(a.size)   # unnormalised idft
    c=np.fft.fftshift(c)
    py,px=np.unravel_index(np.argmax(c),c.shape)
    y0,y1,x0,x1=max(0,py-2),min(h,py+3),max(0,px-2),min(w,px+3)
    win5=c[y0:y1,x0:x1].astype(np.float64)
    ys,xs=np.mgrid[y0:y1,x0:x1]
    s=win5.sum(); cx=(win5*xs).sum()/s; cy=(win5*ys).sum()/s
    return (w/2.0-cx, h/2.0-cy), s/(h*w), (px,py)
print("np ",pc(rb,db,hann))

def pc2(a,b,win,variant):
    A=np.fft.rfft2((a*win).astype(np.float32)); B=np.fft.rfft2((b*win).astype(np.float32))   # half spectrum along x
    P=(A*np.conj(B)).astype(np.complex64)
    m=np.abs(P).astype(np.float32)
    eps=np.float32(np.finfo(np.float32).eps)
    if variant=="std":
        C=P*m/(m*m+eps)
    elif variant=="im=mag":
        Bm=m+1j*m
        C=P*np.conj(Bm)/(np.abs(Bm)**2+eps)
        # purely real entries of the packed format (kx=0 and Nyquist columns are packed along y; treat as std there)
        C[:,0]=(P*m/(m*m+eps))[:,0]; C[:,-1]=(P*m/(m*m+eps))[:,-1]
    c=np.fft.irfft2(C,s=a.shape).astype(np.float32)*(a.size)
    c=np.fft.fftshift(c)
    py,px=np.unravel_index(np.argmax(c),c.shape)
    y0,y1,x0,x1=max(0,py-2),min(h,py+3),max(0,px-2),min(w,px+3)
    win5=c[y0:y1,x0:x1].astype(np.float64)
    ys,xs=np.mgrid[y0:y1,x0:x1]
    s=win5.sum(); cx=(win5*xs).sum()/s; cy=(win5*ys).sum()/s
    return (w/2.0-cx, h/2.0-cy), s/(h*w), (px,py), win5.round(0)
for v in ("std","im=mag"):
    r=pc2(rb,db,hann,v); print(v,r[0],r[1],r[2]); print(r[3])
