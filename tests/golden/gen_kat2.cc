// gen_kat2.cc -- regenerates the inputs of KAT-2 (lite/kernels/cuda/conv_compute_test.cc:300-376):
// mt19937(100) + uniform_real_distribution<double>, x in (-36,36) 3x3x3, w in (-10,10) 4x3x3x3.
// The printed q column reproduces the test's (commented-out) expected vector {0,7,8,1}.
// Output was pasted into tests/golden/kat_cuda_conv_int8.json.  g++ -O1 gen_kat2.cc && ./a.out
#include <random>
#include <cstdio>
#include <cstdint>
#include <cmath>
static std::mt19937 mt(100);
static float random_num(float low, float high){ std::uniform_real_distribution<double> dist(low, high); return dist(mt);}
int main(int argc,char**argv){
  int skip = argc>1? atoi(argv[1]):0;
  for(int i=0;i<skip;i++) random_num(0,1);
  int8_t x[27], w[108];
  for(int i=0;i<27;i++) x[i]=(int8_t)random_num(-36,36);
  for(int i=0;i<108;i++) w[i]=(int8_t)random_num(-10,10);
  printf("x:"); for(int i=0;i<27;i++) printf(" %d",x[i]); printf("\nw:"); for(int i=0;i<108;i++) printf(" %d",w[i]); printf("\n");
  for(int k=0;k<4;k++){ int acc=0; for(int i=0;i<27;i++) acc+=x[i]*w[k*27+i];
    float s = (0.01f*(k+1)*1.0f)/2.0f; float b=(k+1.0f)*(1.f/2.f); float v=fmaf((float)acc,s,b); v = v>0?v:0; 
    printf("k=%d acc=%d v=%f q=%d\n",k,acc,v,(int)nearbyintf(fminf(fmaxf(v,-128),127)));}
}
