#include "../../nonmyopia_b200/csrc/fastmath.cuh"
#include <cstdio>
#include <cstdlib>
#include <cmath>
using namespace nmb;
int main(){
  double me=0,mm=0,ms=0,mc=0,mg=0,mt=0; srand(1);
  for(int i=0;i<4000000;i++){
    double u=(double)rand()/RAND_MAX;
    double x=(u-0.5)*2*60; if(i%4==0) x=(u-0.5)*2e-3; if (i%4==1) x=(u-0.5)*4;
    long double rex=expl((long double)x), rem=expm1l((long double)x);
    double e=nmb_exp(x), m=nmb_expm1(x);
    me=fmax(me,fabs((double)((e-rex)/rex)));
    if (x> -59) mm=fmax(mm,fabs((double)((m-rem)/ (fabsl(rem)>1e-300L?rem:1))));
    double xs=(u-0.5)*2*200; if(i%3==0) xs=(u-0.5)*1e-3;
    double s,c; nmb_sincospi(xs,&s,&c);
    long double rs=sinl(3.14159265358979323846264338327950288L*(long double)xs), rc=cosl(3.14159265358979323846264338327950288L*(long double)xs);
    ms=fmax(ms,fabs((double)(s-rs))); mc=fmax(mc,fabs((double)(c-rc)));
    double xt=(u-0.5)*2*25; if(i%3==0) xt=(u-0.5)*1e-2;
    mg=fmax(mg,fabs(nmb_sigmoid(xt)-(double)(1.0L/(1.0L+expl(-(long double)xt)))));
    long double rt=tanhl((long double)xt);
    mt=fmax(mt,fabs((double)((nmb_tanh(xt)-rt)/(fabsl(rt)>1e-300L?rt:1))));
  }
  printf("exp rel %.3g expm1 rel %.3g sin abs %.3g cos abs %.3g sigmoid abs %.3g tanh rel %.3g\n",me,mm,ms,mc,mg,mt);
  printf("edge: expm1(-100)=%g exp(-800)=%g sigmoid(100)=%.17g tanh(-50)=%.17g expm1(0)=%g sincospi(0.5): ",nmb_expm1(-100),nmb_exp(-800),nmb_sigmoid(100),nmb_tanh(-50),nmb_expm1(0.0));
  double s,c; nmb_sincospi(0.5,&s,&c); printf("%g %g\n",s,c);
}
