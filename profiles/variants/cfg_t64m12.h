#define D4B_FLAT_THREADS {0, 0, 32, 64, 64, 128, 256, 256, 64}
#define D4B_FLAT_MINB {0, 0, 1, 12, 12, 5, 2, 3, 8}
