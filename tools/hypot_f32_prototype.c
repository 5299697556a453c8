/* FP32 prototype of (float)hypot-in-double used to develop an FP32 fast path for the TV-L1 dual update (round 2).
 * The CUDA version was bit-exact and SLOWER than the FP64 formula (profiles/r02_tvl1_hypot_blocks_ab.log), so it is
 * not in the library; this host-side check (gcc -O2 -ffp-contract=off hypot_f32_prototype.c -lm) is kept as the record of
 * the algorithm: approximate root, exact residual from error-free products/sums, correction in ulps, FP64 near ties. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
static inline uint32_t f2u(float f){uint32_t u;memcpy(&u,&f,4);return u;}
static inline float u2f(uint32_t u){float f;memcpy(&f,&u,4);return f;}
static long nslow=0;
static float slow(float a,float b){nslow++;return (float)sqrt((double)a*(double)a+(double)b*(double)b);}
static int perturb=0;
static float hyp(float a,float b){
    const float ah=a*a, bh=b*b;
    const float s=ah+bh;
    if(!(s>=0x1p-60f && s<=0x1p100f)) return (a==0.0f&&b==0.0f)?0.0f:slow(a,b);
    const float al=fmaf(a,a,-ah), bl=fmaf(b,b,-bh);
    const float bb=s-ah;
    const float sl=(ah-(s-bb))+(bh-bb);
    float g0=sqrtf(s);
    g0=u2f(f2u(g0)+perturb);                      // emulate an approximate square root
    const uint32_t gb=f2u(g0), man=gb&0x7fffffu;
    if(man<8u||man>0x7ffff7u) return slow(a,b);
    const float ph=g0*g0, pl=fmaf(g0,g0,-ph);
    const float d=(s-ph)+(((sl+al)+bl)-pl);
    const float u=u2f(gb&0x7f800000u)*0x1p-23f;  // ulp(g0)
    const float q=d*0.5f*(1.0f/(g0*u));
    const float k=rintf(q);
    if(!(fabsf(q-k)<0.5f-0x1p-12f) || !(fabsf(k)<=4.0f)) return slow(a,b);
    return u2f(gb+(int)k);
}
static uint64_t st=88172645463325252ull;
static uint32_t rnd(){st^=st<<13;st^=st>>7;st^=st<<17;return (uint32_t)(st>>16);}
int main(int argc,char**argv){
    long n=argc>1?atol(argv[1]):100000000; long bad=0;
    for(perturb=-2;perturb<=2;perturb++){
        nslow=0; bad=0;
        for(long i=0;i<n;i++){
            uint32_t r=rnd(), r2=rnd(), r3=rnd();
            float a,b;
            int mode=r3%8;
            int ea=97+(r3>>8)%50, eb=mode<3?ea+(int)((r3>>16)%5)-2:97+(r3>>20)%50;
            a=u2f((r&0x807fffffu)|((uint32_t)ea<<23)); b=u2f((r2&0x807fffffu)|((uint32_t)eb<<23));
            if(mode==3) b=0; if(mode==4){a=(float)(int)(r%4096); b=(float)(int)(r2%4096);} if(mode==5) b=a;
            if(mode==6){ a=u2f(f2u(a)&0xffff0000u); b=u2f(f2u(b)&0xffff0000u);}
            float g=hyp(a,b), ref=(float)sqrt((double)a*(double)a+(double)b*(double)b);
            if(f2u(g)!=f2u(ref)){ if(bad<5) printf("BAD a=%a b=%a got %a ref %a\n",a,b,g,ref); bad++; }
        }
        printf("perturb %+d: %ld samples, %ld mismatches, slow path %.4f %%\n",perturb,n,bad,100.0*nslow/n);
    }
    return 0;
}
