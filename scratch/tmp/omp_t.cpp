#include <omp.h>
#include <cstdio>
#include <cmath>
#include <chrono>
int main(){ const int ts[4]={1,2,4,8}; for(int t: ts){ omp_set_num_threads(t); auto t0=std::chrono::steady_clock::now(); double s=0;
#pragma omp parallel for reduction(+:s) schedule(static)
 for(long i=0;i<200000000;i++) s+=std::sqrt((double)i);
 auto t1=std::chrono::steady_clock::now(); printf("%d threads %.1f ms (%g)\n",t,std::chrono::duration<double,std::milli>(t1-t0).count(),s);} }
