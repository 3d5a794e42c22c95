/* Host check of the operation sequence used by nums_b200/csrc/libm_exact.cuh (log1p_glibc_fma) against the C
 * library's log1p, bit for bit:  gcc -O2 -mfma -ffp-contract=off -o /tmp/check_log1p tools/check_log1p_exact.c -lm
 * Prints the number of mismatches over 4*10^8 arguments (0 on glibc 2.39 / x86-64 with FMA). */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
static const double ln2_hi=6.93147180369123816490e-01, ln2_lo=1.90821492927058770002e-10;
static const double Lp1=6.666666666666735130e-01,Lp2=3.999999999940941908e-01,Lp3=2.857142874366239149e-01,Lp4=2.222219843214978396e-01,Lp5=1.818357216161805012e-01,Lp6=1.531383769920937332e-01,Lp7=1.479819860511658591e-01;
static inline int32_t hi(double x){uint64_t u;memcpy(&u,&x,8);return (int32_t)(u>>32);}
static inline double sethi(double x,int32_t h){uint64_t u;memcpy(&u,&x,8);u=(u&0xffffffffull)|((uint64_t)(uint32_t)h<<32);memcpy(&x,&u,8);return x;}
/* glibc 2.39 __log1p_fma (sysdeps/ieee754/dbl-64/s_log1p.c built with -mfma -mavx2), operation by operation */
double log1p_fma(double x){
  double hfsq,f,c=0,s,z,R,u; int32_t k,hx,hu,ax;
  hx=hi(x); ax=hx&0x7fffffff;
  if(hx<0x3FDA827A){
    if(ax>=0x3ff00000){ if(x==-1.0) return -INFINITY; return NAN; }
    if(ax<0x3e200000){ if(ax<0x3c900000) return x; return fma(-(x*x),0.5,x); }
    if(hx>0||hx<=((int32_t)0xbfd2bec3)){ k=0; f=x; hu=1; hfsq=(x*0.5)*x; goto poly; }
  } else if(hx>=0x7ff00000) return x+x;
  if(hx<0x43400000){ u=1.0+x; hu=hi(u); k=(hu>>20)-1023; c=(k>0)?1.0-(u-x):x-(u-1.0); c/=u; }
  else { u=x; hu=hi(u); k=(hu>>20)-1023; c=0; }
  hu&=0x000fffff;
  if(hu<0x6a09e){ u=sethi(u,hu|0x3ff00000); }
  else { k+=1; u=sethi(u,hu|0x3fe00000); hu=(0x00100000-hu)>>2; }
  f=u-1.0; hfsq=(f*0.5)*f;
  if(hu==0){
    if(f==0.0){ if(k==0) return 0.0; return fma((double)k,ln2_hi,fma((double)k,ln2_lo,c)); }
    R=fma(-f,0.66666666666666666,1.0)*hfsq;
    if(k==0) return f-R;
    return fma((double)k,ln2_hi,-((R-fma((double)k,ln2_lo,c))-f));
  }
poly:
  s=f/(f+2.0); z=s*s;
  { double R2=fma(z,Lp3,Lp2), R3=fma(z,Lp5,Lp4), R4=fma(z,Lp7,Lp6), z2=z*z, z4=z2*z2, z6=z2*z4;
    double t=z2*R2; t=fma(z,Lp1,t); t=fma(z4,R3,t); R=fma(z6,R4,t); }
  { double t=(R+hfsq)*s;
    if(k==0) return f-(hfsq-t);
    double c2=fma((double)k,ln2_lo,c);
    return fma((double)k,ln2_hi,-((hfsq-(c2+t))-f)); }
}
int main(){
  uint64_t st=88172645463325252ull; long bad=0,n=400000000;
  for(long i=0;i<n;i++){
    st^=st<<13; st^=st>>7; st^=st<<17;
    double u=(double)(st>>11)*(1.0/9007199254740992.0);
    if(i%7==0) u=u*1e-3; if(i%11==0) u=1.0-u*1e-6; if(i%13==0) u=u*1e-9; if(i%17==0) u=-u*3; if(i%19==0) u=-u*1e6;
    double x=-u, r=log1p(x), a=log1p_fma(x);
    if(memcmp(&r,&a,8) && !(isnan(r)&&isnan(a))){ if(bad<8) printf("x=%a ref=%a got=%a\n",x,r,a); bad++; }
  }
  printf("n=%ld mismatches: %ld\n",n,bad); return 0; }
