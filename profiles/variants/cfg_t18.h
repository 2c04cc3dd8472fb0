#define D4B_CURVED_THREADS {0}
#define D4B_CURVED_MINB {0, 0, 16, 16, 18, 8, 0, 0, 0}
