#include <stdio.h>
#include <stdint.h>
#include <math.h>
#include <string.h>
static uint64_t s = 88172645463325252ULL;
static inline uint64_t rnd(void){ s ^= s<<13; s ^= s>>7; s ^= s<<17; return s; }
static inline double mk(double x, double d, double c){ double q0 = x*c; double r = fma(-d,q0,x); double q = fma(r,c,q0); return copysign(q,x);}
int main(){
  const double c3 = 1.0/3.0, c6 = 1.0/6.0;
  long bad3=0,bad6=0,n=0; int minbad_e3=9999,maxbad_e3=-1,minbad_e6=9999,maxbad_e6=-1;
  for (int pass=0; pass<3; pass++) for (long i=0;i<200000000L;i++){
    uint64_t b = rnd();
    uint64_t e;
    if (pass==0) e = (b>>52)&0x7ff;            // all exponents
    else if (pass==1) e = (b>>52)&0x3f;        // tiny exponents 0..63
    else e = 1000 + ((b>>52)&0x3f);            // mid
    if (e==0x7ff) continue;
    uint64_t bits = (b & 0x800fffffffffffffULL) | (e<<52);
    double x; memcpy(&x,&bits,8);
    double q3 = mk(x,3.0,c3), q6 = mk(x,6.0,c6);
    double t3 = x/3.0, t6 = x/6.0;
    n++;
    if (memcmp(&q3,&t3,8)) { bad3++; if ((int)e<minbad_e3)minbad_e3=e; if((int)e>maxbad_e3)maxbad_e3=e; }
    if (memcmp(&q6,&t6,8)) { bad6++; if ((int)e<minbad_e6)minbad_e6=e; if((int)e>maxbad_e6)maxbad_e6=e; }
  }
  printf("n=%ld bad3=%ld (e %d..%d) bad6=%ld (e %d..%d)\n", n,bad3,minbad_e3,maxbad_e3,bad6,minbad_e6,maxbad_e6);
  double z = -0.0; double q = mk(z,3.0,c3); printf("-0 -> %g signbit %d\n", q, signbit(q)!=0);
  return 0;
}
/* (the integer form used for |x| < 2^-1021 is checked by scripts/check_div6_int.c) */
