"""Adds the debug timeline stamps (compiled only with -DSPS_TIMELINE) to head_bwd_tc.cu.  Idempotent."""
p='scaleprotoseg_b200/csrc/head_bwd_tc.cu'
s=open(p).read()
if "g_tlb" in s:
    raise SystemExit("already instrumented")
s=s.replace('''namespace sps {
namespace tcb {
using namespace sps::tc;
''','''namespace sps {
namespace tcb {
using namespace sps::tc;

#ifdef SPS_TIMELINE
// Debug build only: clock stamps of CTA 0 (role r, event e) -> g_tlb[r * 1024 + e]; read by sps_debug_timeline_bwd.
__device__ long long g_tlb[8 * 1024];
#define TLB(role, ev)                                                                      \\
  do {                                                                                     \\
    if (blockIdx.x == 0 && (ev) < 1024) g_tlb[(role) * 1024 + (ev)] = clock64();            \\
  } while (0)
#else
#define TLB(role, ev) do { } while (0)
#endif
''',1)
lines=s.split('\n')
def ins_after(match, text, nth=0):
    idx=[i for i,l in enumerate(lines) if match in l][nth]
    lines.insert(idx+1, text)
def ins_before(match, text, nth=0):
    idx=[i for i,l in enumerate(lines) if match in l][nth]
    lines.insert(idx, text)
ins_before("if (j == 0) mbar_wait(bar(B_AFULL + ab), (ti >> 1) & 1);", "      if (lane == 0) TLB(0, 4 * u);")
ins_after("mbar_wait(bar(B_SEMPTY + sb), ((u >> 1) & 1) ^ 1);", "      if (lane == 0) TLB(0, 4 * u + 1);")
ins_after("        mma_commit(bar(B_SFULL + sb));", "        TLB(0, 4 * u + 2);")
ins_before("mbar_wait(bar(B_GFULL + gb), (u >> 1) & 1);", "      if (lane == 0) TLB(1, 4 * u);")
ins_after("if (j == 0) mbar_wait(bar(B_DZEMPTY), (ti & 1) ^ 1);", "      if (lane == 0) TLB(1, 4 * u + 1);")
ins_after("        mma_commit(bar(B_GEMPTY + gb));", "        TLB(1, 4 * u + 2);")
ins_before("mbar_wait(bar(B_AEMPTY + ab), ((ti >> 1) & 1) ^ 1);", "      if (m == 0) TLB(2, 4 * ti);")
ins_after("mbar_wait(bar(B_AEMPTY + ab), ((ti >> 1) & 1) ^ 1);", "      if (m == 0) TLB(2, 4 * ti + 1);")
ins_after("      if (lane == 0) mbar_arrive(bar(B_AFULL + ab));", "      if (m == 0) TLB(2, 4 * ti + 2);")
ins_before("mbar_wait(bar(B_SFULL + sb), (u >> 1) & 1);", "        if (threadIdx.x == 0) TLB(3, 8 * u);")
ins_after("mbar_wait(bar(B_SFULL + sb), (u >> 1) & 1);", "        if (threadIdx.x == 0) TLB(3, 8 * u + 1);")
ins_after("        if (lane == 0) mbar_arrive(bar(B_SEMPTY + sb));", "        if (threadIdx.x == 0) TLB(3, 8 * u + 2);")
ins_before("mbar_wait(bar(B_GEMPTY + gb), ((u >> 1) & 1) ^ 1);", "        if (threadIdx.x == 0) TLB(3, 8 * u + 3);")
ins_after("mbar_wait(bar(B_GEMPTY + gb), ((u >> 1) & 1) ^ 1);", "        if (threadIdx.x == 0) TLB(3, 8 * u + 4);")
ins_after("        if (lane == 0) mbar_arrive(bar(B_GFULL + gb));", "        if (threadIdx.x == 0) TLB(3, 8 * u + 5);")
ins_before("mbar_wait(bar(B_DZFULL), ti & 1);", "      if (threadIdx.x == 0) TLB(4, 4 * tj);")
ins_after("mbar_wait(bar(B_DZFULL), ti & 1);", "      if (threadIdx.x == 0) TLB(4, 4 * tj + 1);")
ins_after("      if (lane == 0) mbar_arrive(bar(B_AEMPTY + ab));", "      if (threadIdx.x == 0) TLB(4, 4 * tj + 2);", nth=-1)
s='\n'.join(lines)
s=s.replace("}  // namespace tcb\n","}  // namespace tcb\n\n#ifdef SPS_TIMELINE\nextern \"C\" int sps_debug_timeline_bwd(long long* host, int count) {\n  return (int)cudaMemcpyFromSymbol(host, tcb::g_tlb, sizeof(long long) * (size_t)count);\n}\n#endif\n",1)
open(p,'w').write(s)
