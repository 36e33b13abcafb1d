# oracle/instrument.sed -- TEST INFRASTRUCTURE. Applied by oracle/Makefile to
# /root/reference/founderblockgraph.cpp to generate oracle/_ref/fbg_instrumented.cpp.
# Hook 1: in segment() and segment2elasticValid() (fbg.cpp:664 and :866) the
# locals v, s, prev and boundaries are in scope right after this stderr line.
/std::cerr << "Number of segments: " << boundaries.size() << std::endl;/a\
		FBG_ORACLE_DUMP_NONELASTIC(v, s, prev, boundaries);
# Hook 2: in the elastic DP (fbg.cpp:2023 and its copies) minmaxlength/backtrack
# are in scope right after the "done (optimal segment length" line.
/std::cerr << "done (optimal segment length = " << minmaxlength\[n\] << ")." << std::endl << std::flush;/a\
		FBG_ORACLE_DUMP_ELASTIC(minmaxlength, backtrack);
