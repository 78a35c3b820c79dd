// Host build of the eBayes kernel's Student-t tail (csrc/tdist.cuh): lets a CPU test run the exact GPU arithmetic.
#include "../tdist.cuh"

extern "C" {

// two-sided p-values of n t statistics with df degrees of freedom each (df_stride = 0: one shared df)
void cml_host_t_two_sided(const double* t, const double* df, int df_stride, int n, double* p_out) {
  for (int i = 0; i < n; ++i) p_out[i] = cml::t_two_sided(t[i], df[(long)i * df_stride]);
}

}  // extern "C"
