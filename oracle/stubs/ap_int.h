/* Empty stand-in for Xilinx's <ap_int.h>: device/PublicDefines.h includes it,
 * but the host/test sources compiled into oracle/_ref use only the macros. */
