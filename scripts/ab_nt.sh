for r in 1 2; do
  echo "NT default (512) run $r"; python scripts/dsweep_profile.py c1 device 5 | grep -v '"rep": [01],' | cut -c1-90
  echo "NT 256 run $r"; TCI_DSWEEP_NT=256 python scripts/dsweep_profile.py c1 device 5 | grep -v '"rep": [01],' | cut -c1-90
done
