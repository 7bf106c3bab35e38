#!/bin/sh
# Collects the machine-readable lines of a sweep -- counterpart of scripts/experiments/gather_all.sh:1-20:
#   gather_all.sh <result folder>   ->  aggregate_profiling_data.txt (PROFILING_DATUM lines)
#                                       aggregate_sql_data.txt      (INSERT INTO lines, ready for the reference's tables)
#                                       aggregate_summary.csv       (matrix,kernel,statistic,correct,iteration,trial,time_ms: one line per row)
resultfolder=$1
: > aggregate_profiling_data.txt
: > aggregate_sql_data.txt
echo "matrix,kernel,statistic,correct,iteration,trial,time_ms" > aggregate_summary.csv
find "$resultfolder" -name "result_*.tar.gz" | while read -r z; do (cd "$(dirname "$z")" && tar xzf "$(basename "$z")"); done
find "$resultfolder" -name "*.txt" | sort | while read -r f; do
  grep "PROFILING_DATUM" "$f" >> aggregate_profiling_data.txt
  grep "INSERT INTO" "$f" >> aggregate_sql_data.txt
  # every INSERT carries one or more rows: (<ms>, "<correct>", "<kernel>", <global>, <local>, "<host>", "<device>", "<matrix>",<iteration>,<trial>,"<statistic>", "<experiment>")
  grep "INSERT INTO" "$f" | sed -e 's/.*VALUES *//' | grep -o '([^()]*)' | tr -d '()" ' | \
      awk -F',' '{print $8 "," $3 "," $11 "," $2 "," $9 "," $10 "," $1}' >> aggregate_summary.csv
done
echo "profiling data: $(wc -l < aggregate_profiling_data.txt) lines, sql: $(wc -l < aggregate_sql_data.txt) lines"
