// Developer tool: builds the device plan of a model on the host (no GPU) and prints table sizes and the
// shared-memory layout.  Build: nvcc -DMYO_PLAN_HOST_ONLY -shared -Xcompiler -fPIC -o tools/libplanstats.so tools/plan_stats.cu
#define MYO_PLAN_HOST_ONLY
#include "../myosuite_mjx_b200/csrc/plan.cu"
#include <cstdio>
extern "C" int plan_stats(const myo_model_desc* desc, int nconmax, int njmax) {
  DevModel D; SmemLayout L; std::vector<void*> owned;
  try { myo_build_plan(*desc, nconmax, njmax, D, L, owned); } catch (const std::string& e) { printf("error: %s\n", e.c_str()); return -1; }
  printf("nq %d nv %d nu %d nbody %d nframe %d nlevel %d njnt %d ngeom %d nsite %d nten %d nM %d\n", D.nq, D.nv, D.nu, D.nbody, D.nframe, D.nlevel, D.njnt, D.ngeom, D.nsite, D.nten, D.nM);
  printf("nunit %d nslot %d nentry %d nnz %d nldlop %d ndlevel %d\n", D.nunit, D.nslot, D.nentry, D.nnz, D.nldlop, D.ndlevel);
  printf("ncg %d npair %d neq %d nlimj %d nlimt %d maxbase %d nconmax %d njmax %d\n", D.ncg, D.npair, D.neq, D.nlimj, D.nlimt, D.maxbase, D.nconmax, D.njmax);
  int ngeomunit = 0; for (int u = 0; u < D.nunit; u++) if (D.unit_geom[u] >= 0) ngeomunit++;
  printf("geom units %d\n", ngeomunit);
  int types[8][8] = {{0}}; for (int p = 0; p < D.npair; p++) types[D.cg_type[D.pair_cg1[p]]][D.cg_type[D.pair_cg2[p]]]++;
  for (int a = 0; a < 8; a++) for (int b = 0; b < 8; b++) if (types[a][b]) printf("pairs type %d-%d: %d\n", a, b, types[a][b]);
  printf("smem floats total %d (%d B): persistent %d, overlayA(slots %d ulen %d) overlayB(cgpose %d con %d conJ %d efc %d H %d nvec %d limrow %d basev %d)\n",
         L.total, L.total * 4, L.slots, 6 * D.nslot, 2 * D.nunit, 12 * D.ncg, 12 * D.nconmax, D.maxbase * D.nconmax * D.nv, 10 * D.njmax, D.nv * D.nv, 6 * D.nv, 2 * (D.nlimj + D.nlimt), 2 * D.maxbase * D.nconmax);
  printf("offsets: frames %d cinert %d crb %d cdof %d cdof_dot %d cvel %d cacc %d qM %d qLD %d tenJ %d actlen %d bias %d base %d\n", L.frames, L.cinert, L.crb, L.cdof, L.cdof_dot, L.cvel, L.cacc, L.qM, L.qLD, L.tenJ, L.actlen, L.bias, L.slots);
  return 0;
}
