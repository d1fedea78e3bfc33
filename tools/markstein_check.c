// CPU check of the Markstein quotient used by closures.cuh::div_rn (gcc -O2 -mfma -ffp-contract=off).
// Result on the build machine: N=300000000 plain a*y mismatches=64844270  3-op mismatches=0  5-op mismatches=0
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
static uint64_t s[2]={0x9E3779B97F4A7C15ULL,0xD1B54A32D192ED03ULL};
static inline uint64_t nxt(void){uint64_t a=s[0],b=s[1];s[0]=b;a^=a<<23;s[1]=a^b^(a>>17)^(b>>26);return s[1]+b;}
static inline double rnd_m(int mode){ // mantissa patterns
  uint64_t m=nxt()&((1ULL<<52)-1);
  if(mode==1) m&=~((1ULL<<(nxt()%52))-1);      // trailing zeros
  if(mode==2) m|=((1ULL<<(nxt()%52))-1);       // trailing ones
  if(mode==3) m=(1ULL<<52)-1-(nxt()&0xFF);     // near all ones
  if(mode==4) m=nxt()&0xFF;                    // near power of two
  uint64_t e=1023ULL+(nxt()%40)-20; uint64_t u=(e<<52)|m; double d; memcpy(&d,&u,8); return (nxt()&1)?d:-d;}
int main(int argc,char**argv){
  long N=atol(argv[1]); long bad3=0,bad5=0,badq=0;
  for(long i=0;i<N;i++){
    double a=rnd_m(nxt()%5), b=rnd_m(nxt()%5);
    double y=1.0/b, ref=a/b;
    double q=a*y; double r=fma(-b,q,a); double q1=fma(r,y,q);
    if(q1!=ref) bad3++;
    double r2=fma(-b,q1,a); double q2=fma(r2,y,q1);
    if(q2!=ref) bad5++;
    if(q!=ref) badq++;
  }
  printf("N=%ld plain a*y mismatches=%ld  3-op mismatches=%ld  5-op mismatches=%ld\n",N,badq,bad3,bad5);
  return 0;}
