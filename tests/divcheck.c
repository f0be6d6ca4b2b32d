// Known-answer check for common.cuh div_by_known: with inv = RN(1/b) (b's significand not all ones) the two-FMA quotient
// q = a*inv; r = fma(-b,q,a); q' = fma(r,inv,q) equals the IEEE division a/b bit for bit. Compiled and run by tests/test_oracle.py.
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
static uint64_t s[2]={0x9E3779B97F4A7C15ull,0xD1B54A32D192ED03ull};
static uint64_t nxt(void){uint64_t a=s[0],b=s[1];s[0]=b;a^=a<<23;s[1]=a^b^(a>>17)^(b>>26);return s[1]+b;}
static double mk(int span){uint64_t r=nxt();int e=1023+(int)((r>>52)%(2*span+1))-span;uint64_t b=((uint64_t)e<<52)|(r&0xfffffffffffffull);double d;memcpy(&d,&b,8);return d;}
int main(void){
  long bad=0,n=0;
  for(int rep=0;rep<60;rep++){
    double b = rep<20? mk(3): (rep<40? mk(40): mk(90));
    uint64_t bb; memcpy(&bb,&b,8); if((bb&0xfffffffffffffull)==0xfffffffffffffull) continue;
    double inv=1.0/b;
    for(int i=0;i<200000;i++){
      double a = mk(rep%3==0?2:60);
      if(i%7==0){ /* near multiples of b: hard cases */ double m=(double)(nxt()%100000)+0.5; a=b*m; if(i%14==0) a=nextafter(a, (i&16)?1e300:-1e300);} 
      double q=a*inv; double r=fma(-b,q,a); double q2=fma(r,inv,q);
      if(q2!=a/b) { if(bad<5) printf("mismatch a=%a b=%a got=%a want=%a\n",a,b,q2,a/b); bad++; }
      n++;
    }
  }
  printf("tested %ld mismatches %ld\n",n,bad); return bad!=0;
}
