// Kernel-template instantiation list: (NP = padded active-species count, NSLOT = rows per lane).
#pragma once
#define DFC_DISPATCH_NP(NP_VALUE, CALL)                                  \
    switch (NP_VALUE) {                                                  \
        case 8: CALL(8, 1); break;                                       \
        case 16: CALL(16, 1); break;                                     \
        case 20: CALL(20, 1); break;                                     \
        case 24: CALL(24, 1); break;                                     \
        case 32: CALL(32, 1); break;                                     \
        case 33: CALL(33, 2); break;                                     \
        case 34: CALL(34, 2); break;                                     \
        case 35: CALL(35, 2); break;                                     \
        case 36: CALL(36, 2); break;                                     \
        case 40: CALL(40, 2); break;                                     \
        case 48: CALL(48, 2); break;                                     \
        case 64: CALL(64, 2); break;                                     \
        default: break;                                                  \
    }
