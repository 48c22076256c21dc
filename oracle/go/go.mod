module github.com/intuit/foremast/oracle

go 1.21
