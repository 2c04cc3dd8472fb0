#define D4B_FAST_THREADS {0, 0, 32, 128, 32, 128, 256, 256, 64}
#define D4B_FAST_MINB {0, 0, 16, 5, 20, 4, 1, 1, 5}
