#define D4B_FLAT_THREADS {0, 0, 32, 64, 32, 128, 256, 256, 64}
#define D4B_FLAT_MINB {0, 0, 1, 12, 20, 5, 2, 3, 8}
