// fft_dispatch.cuh — runtime radix -> template dispatch shared by the kernels and the emulator.
#pragma once
#define BLUR_RADIX_SWITCH(r, F)                                                                  \
    switch (r) {                                                                                 \
        case 1: F(1); break;   case 2: F(2); break;   case 3: F(3); break;   case 4: F(4); break;   \
        case 5: F(5); break;   case 6: F(6); break;   case 8: F(8); break;   case 9: F(9); break;   \
        case 10: F(10); break; case 12: F(12); break; case 15: F(15); break; case 16: F(16); break; \
        case 18: F(18); break; case 20: F(20); break; case 24: F(24); break; case 25: F(25); break; \
        case 27: F(27); break; case 30: F(30); break; case 32: F(32); break;                     \
        default: break;                                                                          \
    }
