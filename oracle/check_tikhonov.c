#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <string.h>
static inline float rsq_approx(float x){ // emulate MUFU.RSQ: ~2^-22.4 rel err: perturb exact
    float r = (float)(1.0/sqrt((double)x)); uint32_t u; memcpy(&u,&r,4); u += (rand()%5)-2; memcpy(&r,&u,4); return r; }
static inline float rcp_approx(float x){ float r=(float)(1.0/(double)x); uint32_t u; memcpy(&u,&r,4); u += (rand()%3)-1; memcpy(&r,&u,4); return r; }
#include <string.h>
static float tik2f(float v, float a_hi, float a_lo){
    float p_hi = a_hi*v;
    float p_lo = fmaf(a_hi, v, -p_hi);
    p_lo = fmaf(a_lo, v, p_lo);
    float x_hi = 1.0f + p_hi;
    float bb = x_hi - 1.0f;
    float e = (1.0f - (x_hi - bb)) + (p_hi - bb);
    float x_lo = e + p_lo;
    float r = rsq_approx(x_hi);
    float s_hi = x_hi*r;
    float rho = fmaf(-s_hi, s_hi, x_hi) + x_lo;
    float s_lo = rho*(0.5f*r);
    float d_hi = s_hi + 1.0f;
    float e2 = d_hi - s_hi;
    float d_lo = (1.0f - e2) + s_lo;
    float n = v + v;
    float q0 = rcp_approx(d_hi);
    float y_hi = n*q0;
    float r1 = fmaf(-y_hi, d_hi, n);
    r1 = fmaf(-y_hi, d_lo, r1);
    return fmaf(r1, q0, y_hi);
}
int main(){
    double lams[] = {0.006, 0.0006, 0.06, 1e-5, 0.5, 3.0};
    for (int li=0; li<6; ++li){
        double lam = lams[li]; double a = 2.0*lam; float a_hi=(float)a; float a_lo=(float)(a-(double)a_hi);
        long mism=0, n=0; double maxulp=0;
        for (long i=0;i<20000000;++i){
            double ex = -4.3 + 9.3*rand()/RAND_MAX; float v = (float)pow(10.0, ex);
            float ref = (float)((sqrt(1.0+2.0*lam*(double)v)-1.0)/lam);
            float got = tik2f(v,a_hi,a_lo);
            ++n; if (ref!=got){ ++mism; double u = fabs((double)got-(double)ref)/(nextafterf(ref,INFINITY)-ref); if(u>maxulp)maxulp=u; }
        }
        printf("lambda %g: %ld / %ld differ, max %.2f ulp\n", lam, mism, n, maxulp);
    }
    return 0;
}
