#define D4B_FLAT_THREADS {0, 0, 32, 32, 32, 128, 256, 256, 64}
#define D4B_FLAT_MINB {0, 0, 16, 24, 24, 5, 2, 3, 8}
