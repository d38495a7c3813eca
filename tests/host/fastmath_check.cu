#include "../../nonmyopia_b200/csrc/fastmath.cuh"
#include <cstdio>
#include <cstdlib>
#include <cmath>
using namespace nmb;
static const double tab[NMB_FAST_TAB_LEN] = NMB_FAST_TAB_INIT;
int main(){
  double me=0,mm=0,ms=0,mc=0,mg=0,mt=0; srand(1);
  double te=0,tm=0,ts=0,tc=0,tg=0,tt=0,tl=0;
  const long double PI=3.14159265358979323846264338327950288L;
  for(int i=0;i<4000000;i++){
    double u=(double)rand()/RAND_MAX;
    double x=(u-0.5)*2*60; if(i%4==0) x=(u-0.5)*2e-3; if (i%4==1) x=(u-0.5)*4;
    long double rex=expl((long double)x), rem=expm1l((long double)x);
    double e=nmb_exp(x), m=nmb_expm1(x);
    me=fmax(me,fabs((double)((e-rex)/rex)));
    if (x> -59) mm=fmax(mm,fabs((double)((m-rem)/ (fabsl(rem)>1e-300L?rem:1))));
    e=nmb_exp_tab(x,tab); m=nmb_expm1_tab(x,tab);
    te=fmax(te,fabs((double)((e-rex)/rex)));
    // expm1 through a table entry != 1 carries an ABSOLUTE error of ~1 ulp(1): relative to max(|result|, 2^-7)
    if (x> -59) tm=fmax(tm,fabs((double)((m-rem)/ fmaxl(fabsl(rem),0.0078125L))));
    if (fabs(x)<0.005) tm=fmax(tm,fabs((double)((m-rem)/ (fabsl(rem)>1e-300L?rem:1))));
    double xs=(u-0.5)*2*200; if(i%3==0) xs=(u-0.5)*1e-3;
    double s,c; nmb_sincospi(xs,&s,&c);
    long double rs=sinl(PI*(long double)xs), rc=cosl(PI*(long double)xs);
    ms=fmax(ms,fabs((double)(s-rs))); mc=fmax(mc,fabs((double)(c-rc)));
    nmb_sincospi_tab(xs,tab,&s,&c);
    ts=fmax(ts,fabs((double)(s-rs))); tc=fmax(tc,fabs((double)(c-rc)));
    double xt=(u-0.5)*2*25; if(i%3==0) xt=(u-0.5)*1e-2;
    long double rg=1.0L/(1.0L+expl(-(long double)xt));
    mg=fmax(mg,fabs(nmb_sigmoid(xt)-(double)rg));
    tg=fmax(tg,fabs(nmb_sigmoid_tab(xt,tab)-(double)rg));
    long double rt=tanhl((long double)xt);
    mt=fmax(mt,fabs((double)((nmb_tanh(xt)-rt)/(fabsl(rt)>1e-300L?rt:1))));
    tt=fmax(tt,fabs((double)((nmb_tanh_tab(xt,tab)-rt)/fmaxl(fabsl(rt),0.0078125L))));
    double xe=(u-0.5)*2*30;
    long double rl= xe>0 ? (long double)xe : expm1l((long double)xe);
    tl=fmax(tl,fabs((double)((nmb_elu_tab(xe,tab)-rl)/fmaxl(fabsl(rl),0.0078125L))));
  }
  printf("exp rel %.3g expm1 rel %.3g sin abs %.3g cos abs %.3g sigmoid abs %.3g tanh rel %.3g\n",me,mm,ms,mc,mg,mt);
  printf("tab: exp rel %.3g expm1 rel %.3g sin abs %.3g cos abs %.3g sigmoid abs %.3g tanh rel %.3g elu rel %.3g\n",te,tm,ts,tc,tg,tt,tl);
  printf("edge: expm1(-100)=%g exp(-800)=%g sigmoid(100)=%.17g tanh(-50)=%.17g expm1(0)=%g sincospi(0.5): ",nmb_expm1(-100),nmb_exp(-800),nmb_sigmoid(100),nmb_tanh(-50),nmb_expm1(0.0));
  double s,c; nmb_sincospi(0.5,&s,&c); printf("%g %g\n",s,c);
  printf("tabedge: expm1(-100)=%g exp(-800)=%g sigmoid(100)=%.17g tanh(-50)=%.17g expm1(0)=%g sincospi(0.5): ",nmb_expm1_tab(-100,tab),nmb_exp_tab(-800,tab),nmb_sigmoid_tab(100,tab),nmb_tanh_tab(-50,tab),nmb_expm1_tab(0.0,tab));
  nmb_sincospi_tab(0.5,tab,&s,&c); printf("%g %g", s, c); nmb_sincospi_tab(-1.0,tab,&s,&c); printf(" %g %g\n",s,c);
}
