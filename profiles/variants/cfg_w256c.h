#define D4B_FLAT_THREADS {0, 0, 32, 64, 32, 128, 256, 128, 64}
#define D4B_FLAT_MINB {0, 0, 1, 12, 24, 5, 1, 4, 8}
