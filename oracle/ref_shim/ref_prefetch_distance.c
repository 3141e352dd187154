/* K/csr.cpp:30 declares `extern int prefetch_distance` which BENCH/ never defines (SURVEY.md 8c gotcha 1);
 * the reference links only thanks to -flto dead-code removal.  TEST INFRASTRUCTURE ONLY. */
int prefetch_distance = 1;
