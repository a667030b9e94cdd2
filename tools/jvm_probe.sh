#!/bin/bash
# SURVEY.md section 8(c), last row: is there a JVM or a JPMML jar on the GPU box?  (There is none in the build container.)
# usage (on the GPU box): bash tools/jvm_probe.sh > gpurun_out/jvm_probe.txt 2>&1
echo "## which java javac scala sbt mvn gradle"; for t in java javac scala sbt mvn gradle; do printf '%s: ' $t; command -v $t || echo "not found"; done
echo "## java -version"; java -version 2>&1 || true
echo "## JAVA_HOME=${JAVA_HOME:-unset}"
echo "## ls ~/.m2 ~/.ivy2 ~/.sbt /usr/lib/jvm /opt/java /usr/java"; ls -d ~/.m2 ~/.ivy2 ~/.sbt /usr/lib/jvm /opt/java /usr/java 2>&1
echo "## find / -name 'pmml-evaluator*.jar' -o -name 'pmml-model*.jar' -o -name 'flink-*.jar' -o -name 'libjvm.so'"
find / -xdev \( -name 'pmml-evaluator*.jar' -o -name 'pmml-model*.jar' -o -name 'flink-*.jar' -o -name 'libjvm.so' -o -name 'rt.jar' \) 2>/dev/null | head -20
echo "## (end of probe: empty lists above mean none found)"
