// Test infrastructure: symrelay_plan_bucketed (plan.h) on a BASELINE-config-3-shaped batch in arrival order: every
// descriptor in exactly one tile, tile kinds and block sums legal, perm a permutation that maps sorted to original.
#include <cstdio>
#include <cstring>
#include <cmath>
#include <ctime>
#include <random>
#include <algorithm>
#define SYMRELAY_HAVE_AVX2_SCAN
#include "../../symmetry_b200/csrc/plan.h"
static double now(){timespec a; clock_gettime(CLOCK_MONOTONIC,&a); return a.tv_sec*1e6+a.tv_nsec*1e-3;}
int main(){ const uint32_t n=131072; std::vector<symrelay_desc> d(n), s(n); std::vector<symrelay_tile> t(n); std::vector<uint32_t> perm(n);
 std::mt19937 rng(5); std::uniform_real_distribution<double> U(1e-9,1.0); std::vector<uint32_t> L(n/2);
 for(auto&x:L){ double v=32.0*pow(U(rng),-1/1.2); x=(uint32_t)std::min(16384.0,floor(v)); }
 uint32_t io=0; for(uint32_t i=0;i<n;i++){ bool op=i>=n/2; uint32_t len=L[i%(n/2)]; d[i].slot=i%(n/2); d[i].len_flags=len|(op?SYMRELAY_F_OPEN:0);
   if(!op){ d[i].in_off=io; d[i].out_off=io+12; } else { d[i].in_off=io+12; d[i].out_off=io; } io+=((len+20+12+15)&~15u);} 
 symrelay_plan_check chk; chk.in_bytes=0xfffffff0u; chk.out_bytes=0xfffffff0u; chk.max_streams=1<<20; uint32_t nb=0; int64_t r=0;
 double best=1e9; for(int k=0;k<30;k++){ double a=now(); r=symrelay_plan_bucketed(d.data(),n,t.data(),n,&nb,&chk,s.data(),perm.data()); best=std::min(best,now()-a);} 
 // validate: tiles cover every sorted descriptor exactly once, block sums <= 32, perm is a permutation
 std::vector<int> seen(n,0), pseen(n,0); uint64_t lanes=0; bool ok=true;
 for(int64_t k=0;k<r;k++){ uint32_t f=t[k].first, cnt=t[k].n_kind&0xffff, kind=t[k].n_kind>>16; const bool wide=(kind&4u)!=0; if(wide) kind&=~4u; uint32_t sum=0; for(uint32_t j=0;j<cnt;j++){ seen[f+j]++; uint32_t c=2+(((s[f+j].len_flags&0xffffff)+63)>>6); sum+=c; if(kind==2 && c!=2+(((s[f].len_flags&0xffffff)+63)>>6)) ok=false; if(kind==1 && c<=32) ok=false; if(wide && (kind!=1 || c<128)) ok=false; if(kind!=1 && c>32) ok=false;} if(kind==0 && sum>32) ok=false; if(kind==2 && cnt!=32) ok=false; lanes+=sum; }
 // longest first: along the small-tile list the frames' block counts never grow, thread tiles and packed tiles alike
 // (what starts last in a batch that needs two rounds of warps must be the short tiles), and the same holds inside the
 // strided part and inside the big-frame part of the bulk class
 { uint32_t prev_small=0xffffffffu, prev_str=0xffffffffu, prev_big_class=0xffffffffu;
   for(int64_t k=0;k<r;k++){ const uint32_t kind=(t[k].n_kind>>16)&3u, lanes_per=(t[k].n_kind>>24)&63u, c=2+(((s[t[k].first].len_flags&0xffffff)+63)>>6);
     if(kind!=1){ if(c>prev_small) ok=false; prev_small=c; }
     else if(lanes_per){ if(c>prev_str) ok=false; prev_str=c; }
     else { const uint32_t cls=31-__builtin_clz(c); if(cls>prev_big_class) ok=false; prev_big_class=cls; } } }
 for(uint32_t i=0;i<n;i++){ if(seen[i]!=1) ok=false; pseen[perm[i]]++; if(memcmp(&s[i],&d[perm[i]],16)) ok=false; }
 for(uint32_t i=0;i<n;i++) if(pseen[i]!=1) ok=false;
 uint32_t nb2; int64_t r2=symrelay_plan_inline(d.data(),n,t.data(),n,&nb2,&chk);
 // the plan as one number (the test runs this program with and without the planner's worker threads: same plan), and the
 // same call many more times, to give a race between the caller and the workers a chance to show
 auto fnv=[&](){ uint64_t h=1469598103934665603ull; auto mix=[&](const void*p,size_t nbytes){ const uint8_t*b=(const uint8_t*)p; for(size_t i=0;i<nbytes;i++){ h^=b[i]; h*=1099511628211ull; } };
   mix(t.data(),(size_t)r*sizeof(symrelay_tile)); mix(s.data(),(size_t)n*16); mix(perm.data(),(size_t)n*4); return h; };
 r=symrelay_plan_bucketed(d.data(),n,t.data(),n,&nb,&chk,s.data(),perm.data());
 const uint64_t h0=fnv();
 for(int k=0;k<300;k++){ std::fill(perm.begin(),perm.end(),0u); int64_t rr=symrelay_plan_bucketed(d.data(),n,t.data(),n,&nb,&chk,s.data(),perm.data()); if(rr!=r||fnv()!=h0){ ok=false; break; } }
 printf("bucketed: %lld tiles (%u bulk) in %.0f us, valid=%d, plan=%016llx; greedy in arrival order: %lld tiles\n",(long long)r,nb,best,(int)ok,(unsigned long long)h0,(long long)r2); return ok?0:1; }
