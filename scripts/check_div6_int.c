/* div6_int of fast_common.cuh against true division for every class of tiny input. */
#include <stdio.h>
#include <stdint.h>
#include <string.h>
static uint64_t s = 88172645463325252ULL;
static inline uint64_t rnd(void){ s ^= s<<13; s ^= s>>7; s ^= s<<17; return s; }
static double div6_int(double x){ uint64_t b; memcpy(&b,&x,8); uint64_t k=b&0x7fffffffffffffffULL; uint64_t q=k/6; uint64_t rem=k-6*q;
  q += (rem>3 || (rem==3 && (q&1))) ? 1:0; uint64_t o=(b&0x8000000000000000ULL)|q; double r; memcpy(&r,&o,8); return r; }
int main(){ long bad=0,n=0; for(long i=0;i<400000000L;i++){ uint64_t b=rnd(); uint64_t bits=(b&0x801fffffffffffffULL); /* e in {0,1} */
  double x; memcpy(&x,&bits,8); double t=x/6.0, q=div6_int(x); n++; if(memcmp(&t,&q,8)) bad++; }
  for (uint64_t k=0;k<2000000;k++){ double x; memcpy(&x,&k,8); double t=x/6.0,q=div6_int(x); n++; if(memcmp(&t,&q,8)) bad++; }
  printf("n=%ld bad=%ld\n",n,bad); return bad!=0; }
