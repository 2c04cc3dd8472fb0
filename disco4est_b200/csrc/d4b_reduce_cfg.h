// d4b_reduce_cfg.h -- launch geometry of the two-pass dot product, shared by d4b_vec_dot (d4b_solver.cu)
// and d4b_vec_dot_noalloc (d4b_static.cu): the same grid and the same trees give bit-identical sums.
#ifndef D4B_REDUCE_CFG_H
#define D4B_REDUCE_CFG_H
constexpr int kD4bRedBlocks = 592;    // 4 x 148 SMs
constexpr int kD4bRedThreads = 256;
#endif
